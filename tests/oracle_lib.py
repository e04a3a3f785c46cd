"""ctypes binding of the CPU oracle (oracle/liblbm_oracle*.so) plus input-file helpers.

TEST INFRASTRUCTURE: imported only by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product never imports this module.
"""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

_f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_c_int, _c_float = ctypes.c_int, ctypes.c_float


def build_oracle():
    """(Re)build the oracle shared libraries with gcc (seconds)."""
    subprocess.run(["make", "-C", ORACLE_DIR, "liblbm_oracle.so", "liblbm_oracle_omp.so"],
                   check=True, stdout=subprocess.DEVNULL)


def load(omp=False):
    name = "liblbm_oracle_omp.so" if omp else "liblbm_oracle.so"
    path = os.path.join(ORACLE_DIR, name)
    if not os.path.exists(path):
        build_oracle()
    lib = ctypes.CDLL(path)
    lib.oracle_init.argtypes = [_c_int, _c_int, _c_float, _f32p]
    lib.oracle_init.restype = None
    lib.oracle_accelerate_flow.argtypes = [_c_int, _c_int, _c_float, _c_float, _f32p, _i32p]
    lib.oracle_accelerate_flow.restype = None
    lib.oracle_run_ranks.argtypes = [_c_int, _c_int, _c_int, _c_float, _c_float, _c_float,
                                     _i32p, _f32p, _f32p, _c_int, _c_int]
    lib.oracle_run_ranks.restype = _c_int
    lib.oracle_set_exact_sums.argtypes = [ctypes.c_void_p]
    lib.oracle_set_exact_sums.restype = None
    lib.oracle_run.argtypes = [_c_int, _c_int, _c_int, _c_float, _c_float, _c_float,
                               _i32p, _f32p, _f32p, _c_int]
    lib.oracle_run.restype = _c_int
    lib.oracle_timestep.argtypes = [_c_int, _c_int, _c_float, _i32p, _f32p, _f32p,
                                    ctypes.POINTER(_c_float)]
    lib.oracle_timestep.restype = _c_int
    lib.oracle_scale_av_vels.argtypes = [_f32p, _c_int, _c_int, _c_int, _c_int]
    lib.oracle_scale_av_vels.restype = None
    lib.oracle_final_state.argtypes = [_c_int, _c_int, _c_float, _i32p, _f32p,
                                       _f32p, _f32p, _f32p, _f32p]
    lib.oracle_final_state.restype = None
    lib.oracle_av_velocity.argtypes = [_c_int, _c_int, _i32p, _f32p]
    lib.oracle_av_velocity.restype = _c_float
    lib.oracle_reynolds.argtypes = [_c_int, _c_int, _c_int, _c_float, _i32p, _f32p]
    lib.oracle_reynolds.restype = _c_float
    lib.oracle_total_density.argtypes = [_c_int, _c_int, _f32p]
    lib.oracle_total_density.restype = _c_float
    return lib


class Oracle:
    """Thin numpy front end; lattices are float32 arrays of shape (9, ny, nx)."""

    def __init__(self, omp=False):
        self.lib = load(omp)

    def init(self, nx, ny, density):
        f = np.empty((9, ny, nx), np.float32)
        self.lib.oracle_init(nx, ny, density, f)
        return f

    def run(self, nx, ny, iters, density, accel, omega, obst, f=None, ranks=1, exact=False):
        """Returns (lattice, tot_u[iters] un-normalised); with exact=True additionally the per-step sums of the
        same per-cell fp32 speeds accumulated in double (float64[iters])."""
        from_state = f is not None
        f = np.ascontiguousarray(f, np.float32).copy() if from_state else np.empty((9, ny, nx), np.float32)
        tot_u = np.zeros(max(iters, 1), np.float32)
        sums = np.zeros(max(iters, 1), np.float64)
        self.lib.oracle_set_exact_sums(sums.ctypes.data if exact else None)
        try:
            rc = self.lib.oracle_run_ranks(nx, ny, iters, density, accel, omega,
                                           np.ascontiguousarray(obst, np.int32), f, tot_u, ranks, int(from_state))
        finally:
            self.lib.oracle_set_exact_sums(None)
        if rc != 0:
            raise RuntimeError("oracle_run_ranks failed")
        return (f, tot_u[:iters], sums[:iters]) if exact else (f, tot_u[:iters])

    def accelerate_flow(self, f, density, accel, obst):
        _, ny, nx = f.shape
        g = np.ascontiguousarray(f, np.float32).copy()
        self.lib.oracle_accelerate_flow(nx, ny, density, accel, g, np.ascontiguousarray(obst, np.int32))
        return g

    def timestep(self, f, omega, obst):
        _, ny, nx = f.shape
        dst = np.empty_like(f)
        t = _c_float(0)
        rc = self.lib.oracle_timestep(nx, ny, omega, np.ascontiguousarray(obst, np.int32),
                                      np.ascontiguousarray(f, np.float32), dst, ctypes.byref(t))
        if rc != 0:
            raise RuntimeError("oracle_timestep failed")
        return dst, np.float32(t.value)

    def scale_av_vels(self, tot_u, nx, ny, obstacle_lines):
        av = np.ascontiguousarray(tot_u, np.float32).copy()
        self.lib.oracle_scale_av_vels(av, av.size, nx, ny, obstacle_lines)
        return av

    def final_state(self, f, density, obst):
        _, ny, nx = f.shape
        out = [np.empty((ny, nx), np.float32) for _ in range(4)]
        self.lib.oracle_final_state(nx, ny, density, np.ascontiguousarray(obst, np.int32),
                                    np.ascontiguousarray(f, np.float32), *out)
        return tuple(out)  # ux, uy, u, pressure

    def reynolds(self, f, reynolds_dim, omega, obst):
        _, ny, nx = f.shape
        return np.float32(self.lib.oracle_reynolds(nx, ny, reynolds_dim, omega,
                                                   np.ascontiguousarray(obst, np.int32),
                                                   np.ascontiguousarray(f, np.float32)))

    def total_density(self, f):
        _, ny, nx = f.shape
        return np.float32(self.lib.oracle_total_density(nx, ny, np.ascontiguousarray(f, np.float32)))


# ---------------------------------------------------------------- input files
def read_params(path):
    """The 7-line params file (d2q9-bgk.c:522-528)."""
    with open(path) as fh:
        tok = fh.read().split()
    return dict(nx=int(tok[0]), ny=int(tok[1]), max_iters=int(tok[2]), reynolds_dim=int(tok[3]),
                density=float(np.float32(tok[4])), accel=float(np.float32(tok[5])),
                omega=float(np.float32(tok[6])))


def read_obstacles(path, nx, ny):
    """Returns (int32 mask[ny, nx], number of lines) (d2q9-bgk.c:628-641)."""
    data = np.loadtxt(path, dtype=np.int64, ndmin=2)
    mask = np.zeros((ny, nx), np.int32)
    mask[data[:, 1], data[:, 0]] = 1
    return mask, int(data.shape[0])


def grid_files(grid):
    d = os.path.join(ROOT, "inputs")
    return os.path.join(d, f"input_{grid}.params"), os.path.join(d, f"obstacles_{grid}.dat")


def format_final_state(ux, uy, u, pressure, obst):
    """final_state.dat text exactly as write_values prints it (d2q9-bgk.c:739)."""
    ny, nx = obst.shape
    lines = []
    for jj in range(ny):
        for ii in range(nx):
            lines.append("%d %d %.12E %.12E %.12E %.12E %d\n" % (
                ii, jj, ux[jj, ii], uy[jj, ii], u[jj, ii], pressure[jj, ii], obst[jj, ii]))
    return "".join(lines)


def format_av_vels(av):
    """av_vels.dat text (d2q9-bgk.c:747)."""
    return "".join("%d:\t%.12E\n" % (i, v) for i, v in enumerate(av))

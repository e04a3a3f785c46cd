"""lattice-boltzmann_b200: B200-native D2Q9-BGK timestep engine.

Python-side mirror of the C ABI in include/lbm_b200.h (ctypes; numpy buffers in, numpy buffers
out).  The names follow the reference's own (d2q9-bgk.c): params (t_param, :66-87), speeds
(t_speed, :90-100), accelerate_flow + timestep (:235-339) = `run`, collate (:395-447) =
`av_vels` / `cells`, write_values (:698-752) = `final_state` / `write_values`.

There is no CPU fallback: importing works anywhere (so that symbol/ABI tests can run on a
CPU-only box) but constructing `Lattice` without a usable B200 raises `LbmError`.

The directory name contains a hyphen, so import it with
    importlib.import_module("lattice-boltzmann_b200")
"""
import ctypes
import os

import numpy as np

NSPEEDS = 9
IPC_BLOB_BYTES = 256
_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liblbm_b200.so")


class LbmError(RuntimeError):
    pass


class Params(ctypes.Structure):
    """lbm_params: the hot-path fields of the reference's t_param (d2q9-bgk.c:66-87)."""
    _fields_ = [("nx", ctypes.c_int), ("ny", ctypes.c_int), ("max_iters", ctypes.c_int),
                ("density", ctypes.c_float), ("accel", ctypes.c_float), ("omega", ctypes.c_float),
                ("obstacle_lines", ctypes.c_int)]


class Options(ctypes.Structure):
    _fields_ = [("strict_fp", ctypes.c_int), ("ctas_per_sm", ctypes.c_int), ("threads_per_cta", ctypes.c_int),
                ("engine", ctypes.c_int), ("first_device", ctypes.c_int)]


_lib = None

# name -> (restype, argtypes); every symbol include/lbm_b200.h declares
_P = ctypes.c_void_p
_FP = ctypes.POINTER(ctypes.c_float)
_IP = ctypes.POINTER(ctypes.c_int)
SYMBOLS = {
    "lbm_create": (ctypes.c_int, [ctypes.POINTER(_P), ctypes.POINTER(Params), _P, ctypes.c_int, ctypes.POINTER(Options)]),
    "lbm_destroy": (None, [_P]),
    "lbm_trim_cache": (ctypes.c_int, []),
    "lbm_run": (ctypes.c_int, [_P, ctypes.c_int]),
    "lbm_sync": (ctypes.c_int, [_P]),
    "lbm_get_av_vels": (ctypes.c_int, [_P, _P, ctypes.c_int]),
    "lbm_get_tot_u": (ctypes.c_int, [_P, _P, ctypes.c_int, ctypes.c_int]),
    "lbm_get_cells": (ctypes.c_int, [_P, ctypes.POINTER(_P)]),
    "lbm_set_cells": (ctypes.c_int, [_P, ctypes.POINTER(_P)]),
    "lbm_get_final_state": (ctypes.c_int, [_P, _P, _P, _P, _P]),
    "lbm_av_velocity": (ctypes.c_int, [_P, _FP]),
    "lbm_total_density": (ctypes.c_int, [_P, _FP]),
    "lbm_diagnostics_noalloc": (ctypes.c_int, [_P, _FP, _FP]),
    "lbm_abi_version": (ctypes.c_int, []),
    "lbm_checked_build": (ctypes.c_int, []),
    "lbm_last_error": (ctypes.c_char_p, []),
    "lbm_device_count": (ctypes.c_int, []),
    "lbm_steps_done": (ctypes.c_int, [_P]),
    "lbm_num_slabs": (ctypes.c_int, [_P]),
    "lbm_slab_rows": (ctypes.c_int, [_P, ctypes.c_int, _IP, _IP]),
    "lbm_last_run_ms": (ctypes.c_int, [_P, _FP]),
    "lbm_last_run_launches": (ctypes.c_int, [_P]),
    "lbm_transfer_bytes": (ctypes.c_int, [_P, ctypes.POINTER(ctypes.c_longlong), ctypes.POINTER(ctypes.c_longlong)]),
    "lbm_describe": (ctypes.c_char_p, [_P]),
    "lbm_get_profile": (ctypes.c_int, [_P, ctypes.c_int, _P, ctypes.c_int, _IP, _IP]),
    "lbm_allocate_rows": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_int, _IP, _IP]),
    "lbm_plan_selftest": (ctypes.c_int, [ctypes.c_int] * 4),
    "lbm_host_selftest": (ctypes.c_int, [ctypes.c_int] * 3 + [ctypes.c_longlong]),
    "lbm_selftest_math": (ctypes.c_int, [ctypes.c_longlong, ctypes.c_longlong, ctypes.POINTER(ctypes.c_longlong)]),
    "lbm_selftest_records": (ctypes.c_int, [ctypes.c_int] * 4 + [ctypes.POINTER(ctypes.c_longlong)] * 2),
    "lbm_create_rank": (ctypes.c_int, [ctypes.POINTER(_P), ctypes.POINTER(Params), _P, ctypes.c_int, ctypes.c_int,
                                       ctypes.c_int, ctypes.POINTER(Options)]),
    "lbm_rank_export": (ctypes.c_int, [_P, _P]),
    "lbm_rank_connect": (ctypes.c_int, [_P, _P, _P]),
    "lbm_rank_disconnect": (ctypes.c_int, [_P]),
    "lbm_inv_free_cells": (ctypes.c_float, [_P]),
}


def load_library(path=None):
    """Loads liblbm_b200.so (built in-tree by `make` / __graft_entry__.build()). Fails loudly."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    path = path or os.environ.get("LBM_LIBRARY") or LIB_PATH      # LBM_LIBRARY: e.g. the self-checking build (`make checked`)
    if not os.path.exists(path):
        raise LbmError(f"{path} not found: build it with `make` (nvcc, sm_100a); there is no fallback")
    lib = ctypes.CDLL(path)
    for name, (res, args) in SYMBOLS.items():
        if os.environ.get("LBM_LENIENT_SYMBOLS") and not hasattr(lib, name):
            continue              # dev tool only (tools/ab_variants.sh runs older builds of the library side by side)
        fn = getattr(lib, name)   # AttributeError if the library lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def _err(lib):
    return (lib.lbm_last_error() or b"").decode()


def allocate_rows(ny, size, rank):
    """The reference's row decomposition (allocate_rows, d2q9-bgk.c:484-496): (row_start, rows)."""
    lib = load_library()
    a, b = ctypes.c_int(), ctypes.c_int()
    if lib.lbm_allocate_rows(ny, size, rank, ctypes.byref(a), ctypes.byref(b)):
        raise LbmError(_err(lib))
    return a.value, b.value


def selftest_math(n=1 << 26, seed=1):
    """Mismatches (must be 0) between the kernel's exact div/sqrt fast paths and the IEEE intrinsics."""
    lib = load_library()
    bad = ctypes.c_longlong(-1)
    if lib.lbm_selftest_math(n, seed, ctypes.byref(bad)):
        raise LbmError(_err(lib))
    return bad.value


def selftest_records(writer_device=0, reader_device=0, sweeps=2000, mode=0):
    """(torn, changes): torn 16-byte records seen (must be 0) and value changes observed between consecutive
    looks (> 0: reads and writes overlapped) while one kernel rewrites tagged records and another reads them."""
    lib = load_library()
    torn, changes = ctypes.c_longlong(-1), ctypes.c_longlong(-1)
    if lib.lbm_selftest_records(writer_device, reader_device, sweeps, mode, ctypes.byref(torn), ctypes.byref(changes)):
        raise LbmError(_err(lib))
    return torn.value, changes.value


def trim_cache():
    """Returns the device buffers cached from destroyed lattices to the driver."""
    load_library().lbm_trim_cache()


def device_count():
    return load_library().lbm_device_count()


def _planes(arr):
    ptrs = (_P * NSPEEDS)()
    for k in range(NSPEEDS):
        ptrs[k] = arr[k].ctypes.data
    return ptrs


class Lattice:
    """One simulation: the state initialise() builds plus the time loop of main()."""

    def __init__(self, nx, ny, max_iters, density, accel, omega, obstacles, obstacle_lines=-1, ngpus=1,
                 strict_fp=1, engine=0, ctas_per_sm=0, threads_per_cta=0, first_device=0,
                 rank=None, size=None, device=None):
        self.lib = load_library()
        self.nx, self.ny, self.max_iters = int(nx), int(ny), int(max_iters)
        self.density = float(density)
        obst = np.ascontiguousarray(obstacles, dtype=np.int32).reshape(ny, nx)
        self.obstacles = obst
        self._prm = Params(nx, ny, max_iters, density, accel, omega, obstacle_lines)
        self._opt = Options(strict_fp, ctas_per_sm, threads_per_cta, engine, first_device)
        self._ctx = _P()
        self.rank_mode = rank is not None
        if self.rank_mode:
            rc = self.lib.lbm_create_rank(ctypes.byref(self._ctx), ctypes.byref(self._prm), obst.ctypes.data,
                                          rank, size, device if device is not None else rank, ctypes.byref(self._opt))
            self.row_start, self.rows = allocate_rows(ny, size, rank) if not rc else (0, 0)
        else:
            rc = self.lib.lbm_create(ctypes.byref(self._ctx), ctypes.byref(self._prm), obst.ctypes.data, ngpus,
                                     ctypes.byref(self._opt))
            self.row_start, self.rows = 0, ny
        if rc:
            self._ctx = _P()
            raise LbmError(_err(self.lib))

    # -- life cycle
    def close(self):
        if getattr(self, "_ctx", None):
            self.lib.lbm_destroy(self._ctx)
            self._ctx = _P()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc):
        if rc:
            raise LbmError(_err(self.lib))

    # -- hot path
    def run(self, iters):
        """`iters` trips of the reference's time loop (d2q9-bgk.c:191-197). Asynchronous."""
        self._ck(self.lib.lbm_run(self._ctx, int(iters)))

    def sync(self):
        self._ck(self.lib.lbm_sync(self._ctx))

    # -- results
    @property
    def steps_done(self):
        return self.lib.lbm_steps_done(self._ctx)

    def av_vels(self, n=None):
        n = self.steps_done if n is None else n
        out = np.empty(n, np.float32)
        self._ck(self.lib.lbm_get_av_vels(self._ctx, out.ctypes.data, n))
        return out

    def tot_u(self, first=0, n=None):
        """Un-normalised per-step sums of |u| (timestep()'s return value, d2q9-bgk.c:193) for steps [first, first+n)."""
        n = self.steps_done - first if n is None else n
        out = np.empty(n, np.float32)
        self._ck(self.lib.lbm_get_tot_u(self._ctx, out.ctypes.data, first, n))
        return out

    def cells(self, out=None):
        """The nine planes of this context's rows (collate's cells_global, d2q9-bgk.c:397-443); `out` = a
        caller-owned float32 array (9, rows, nx) to fill instead of a fresh one (e.g. long-lived pinned memory)."""
        f = np.empty((NSPEEDS, self.rows, self.nx), np.float32) if out is None else out
        assert f.shape == (NSPEEDS, self.rows, self.nx) and f.dtype == np.float32 and f.flags.c_contiguous
        self._ck(self.lib.lbm_get_cells(self._ctx, _planes(f)))
        return f

    def set_cells(self, f):
        f = np.ascontiguousarray(f, np.float32)
        assert f.shape == (NSPEEDS, self.ny, self.nx)
        self._ck(self.lib.lbm_set_cells(self._ctx, _planes(f)))

    def final_state(self):
        """(u_x, u_y, |u|, pressure) as write_values computes them (d2q9-bgk.c:710-741)."""
        out = [np.empty((self.rows, self.nx), np.float32) for _ in range(4)]
        self._ck(self.lib.lbm_get_final_state(self._ctx, *[o.ctypes.data for o in out]))
        return tuple(out)

    def av_velocity(self):
        v = ctypes.c_float()
        self._ck(self.lib.lbm_av_velocity(self._ctx, ctypes.byref(v)))
        return np.float32(v.value)

    def reynolds(self, reynolds_dim, omega):
        """calc_reynolds (d2q9-bgk.c:675-678), fp32 like the reference."""
        viscosity = np.float32(1.0) / np.float32(6.0) * (np.float32(2.0) / np.float32(omega) - np.float32(1.0))
        return np.float32(self.av_velocity() * np.float32(reynolds_dim) / viscosity)

    def reynolds_reference_order(self, reynolds_dim, omega):
        """calc_reynolds with av_velocity summed as the reference does (sequential fp32 over the
        unblocked cells, jj outer / ii inner, d2q9-bgk.c:454-480) from the device-computed |u| plane:
        reproduces the reference's printed Reynolds number digit for digit."""
        _, _, u, _ = self.final_state()
        free = u[self.obstacles[self.row_start:self.row_start + self.rows] == 0]
        tot_u = np.cumsum(free, dtype=np.float32)[-1]          # cumsum is a sequential fp32 sum
        av = np.float32(tot_u / np.float32(free.size))
        viscosity = np.float32(1.0) / np.float32(6.0) * (np.float32(2.0) / np.float32(omega) - np.float32(1.0))
        return np.float32(av * np.float32(reynolds_dim) / viscosity)

    def total_density(self):
        v = ctypes.c_float()
        self._ck(self.lib.lbm_total_density(self._ctx, ctypes.byref(v)))
        return np.float32(v.value)

    def diagnostics_noalloc(self):
        """(av_velocity, total_density) through the entry point that allocates nothing (fixed 256 KB device scratch,
        at most LBM_NOALLOC_MAX_ROWS = 16384 rows per GPU)."""
        a, b = ctypes.c_float(), ctypes.c_float()
        self._ck(self.lib.lbm_diagnostics_noalloc(self._ctx, ctypes.byref(a), ctypes.byref(b)))
        return np.float32(a.value), np.float32(b.value)

    # -- introspection
    def last_run_ms(self):
        v = ctypes.c_float()
        self._ck(self.lib.lbm_last_run_ms(self._ctx, ctypes.byref(v)))
        return v.value

    def last_run_launches(self):
        return self.lib.lbm_last_run_launches(self._ctx)

    def transfer_bytes(self):
        """(host->device, device->host) bytes copied through this context so far."""
        a, b = ctypes.c_longlong(), ctypes.c_longlong()
        self._ck(self.lib.lbm_transfer_bytes(self._ctx, ctypes.byref(a), ctypes.byref(b)))
        return a.value, b.value

    def describe(self):
        return (self.lib.lbm_describe(self._ctx) or b"").decode()

    def profile(self, slab=0):
        """[cta, warp, (wait, barrier, publish, total)] SM cycles of the last run (needs LBM_PROFILE=1)."""
        g, w = ctypes.c_int(), ctypes.c_int()
        self._ck(self.lib.lbm_get_profile(self._ctx, slab, None, 0, ctypes.byref(g), ctypes.byref(w)))
        out = np.zeros((g.value, w.value, 4), np.int64)
        self._ck(self.lib.lbm_get_profile(self._ctx, slab, out.ctypes.data, out.size, None, None))
        return out

    def num_slabs(self):
        return self.lib.lbm_num_slabs(self._ctx)

    def inv_free_cells(self):
        return np.float32(self.lib.lbm_inv_free_cells(self._ctx))

    # -- one process per GPU
    def rank_export(self):
        blob = ctypes.create_string_buffer(IPC_BLOB_BYTES)
        self._ck(self.lib.lbm_rank_export(self._ctx, blob))
        return blob.raw

    def rank_connect(self, blob_up, blob_down):
        self._ck(self.lib.lbm_rank_connect(self._ctx, blob_up, blob_down))

    def rank_disconnect(self):
        """Unmaps the neighbours' halo slots; every rank calls it, then a host barrier, then close()."""
        self._ck(self.lib.lbm_rank_disconnect(self._ctx))


def read_params(path):
    """The 7-line params file (d2q9-bgk.c:522-528)."""
    with open(path) as fh:
        tok = fh.read().split()
    if len(tok) < 7:
        raise LbmError(f"could not read param file: {path}")
    return dict(nx=int(tok[0]), ny=int(tok[1]), max_iters=int(tok[2]), reynolds_dim=int(tok[3]),
                density=float(np.float32(tok[4])), accel=float(np.float32(tok[5])), omega=float(np.float32(tok[6])))


def read_obstacles(path, nx, ny):
    """Obstacle list "x y 1" (d2q9-bgk.c:628-641) -> (int32 mask[ny, nx], number of lines)."""
    data = np.loadtxt(path, dtype=np.int64, ndmin=2)
    if data.shape[1] != 3:
        raise LbmError("expected 3 values per line in obstacle file")
    if (data[:, 0] < 0).any() or (data[:, 0] > nx - 1).any():
        raise LbmError("obstacle x-coord out of range")
    if (data[:, 1] < 0).any() or (data[:, 1] > ny - 1).any():
        raise LbmError("obstacle y-coord out of range")
    if (data[:, 2] != 1).any():
        raise LbmError("obstacle blocked value should be 1")
    mask = np.zeros((ny, nx), np.int32)
    mask[data[:, 1], data[:, 0]] = 1
    return mask, int(data.shape[0])


def from_files(paramfile, obstaclefile, **kw):
    """`./d2q9-bgk <paramfile> <obstaclefile>` as an object: returns (Lattice, params dict)."""
    p = read_params(paramfile)
    mask, lines = read_obstacles(obstaclefile, p["nx"], p["ny"])
    lat = Lattice(p["nx"], p["ny"], p["max_iters"], p["density"], p["accel"], p["omega"], mask,
                  obstacle_lines=lines, **kw)
    return lat, p

#!/usr/bin/env python3
"""bench.py - GLUPS of the D2Q9-BGK timestep hot path on N B200s, and % of HBM roofline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    torchrun ... bench.py --gpus N ...            (one rank per GPU, N > 1)

A "step" is one complete run of the workload's time loop (maxIters timesteps of
accelerate_flow + propagate + rebound + collision + av_velocity) from device-resident state.
Workload at every N: 16384x1024 (BASELINE.json's 1/2/4/8-GPU strong-scaling configuration; 1.2 GB
A/B working set, i.e. the honest HBM-bound case), row-decomposed over the N GPUs with
three-population one-row halos pushed over NVLink peer memory from inside the kernel.  One
workload for all N keeps the driver's scaling efficiency meaningful.  At N = 1 the line also
carries BASELINE.json's single-GPU roofline headline, 1024x1024, under "also" (75 MB A/B set:
L2-resident inside a step, so it can exceed the HBM roofline); at N = 2 and N = 4 it carries the
weak-scaling configurations 4096x1024 and 8192x1024 the same way.
value   = lattice updates of all ranks / device time (CUDA events on the launching stream, max
          over ranks), in GLUPS.
e2e     = the same metric through the public C ABI with HOST buffers: lbm_create (mask +
          schedule upload) + lbm_run + lbm_get_av_vels + lbm_get_cells (downloads) + destroy.
roofline= algorithmic bytes (72 B per cell update) / kernel time vs MEASURED_PEAKS.json's HBM figure.
cpu_baseline / --impl reference = the UNMODIFIED reference's OpenMP build (oracle/_ref/ref-omp),
          timed on this box's host cores on a bounded sample (reduced maxIters; LBM cost per
          timestep is constant) from its own "Elapsed Compute time" line.
"""
import argparse
import importlib
import json
import os
import re
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BYTES_PER_UPDATE = 72.0          # 9 fp32 loads + 9 fp32 stores
FALLBACK_HBM_GBS = 6650.0        # /opt/skills/guides/B200_PROFILING.md, only if MEASURED_PEAKS.json is absent
REF_OMP = os.path.join(ROOT, "oracle", "_ref", "ref-omp")


def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


def measured_traffic(grid, updates_per_launch):
    """DRAM bytes of one timestep-kernel launch, from the committed ncu capture (bytes per cell update
    measured on a shorter launch of the same kernel and grid, scaled to this launch's update count)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
            per = json.load(fh)["dram_bytes_per_cell_update"][grid]
        return per * updates_per_launch
    except Exception:
        return None


def grid_files(grid):
    d = os.path.join(ROOT, "inputs")
    return os.path.join(d, f"input_{grid}.params"), os.path.join(d, f"obstacles_{grid}.dat")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index=0):
        self.rows, self.proc, self.idx = [], None, device_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm = sorted(float(r[1]) for r in self.rows if len(r) >= 8 and r[1].replace(".", "").isdigit())
        mx = [float(r[2]) for r in self.rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 8 for i in range(4) if r[4 + i].lower() == "active"})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


# --------------------------------------------------------------------------- CPU reference arm
def run_reference_sample(grid, iters, threads):
    """Runs the unmodified reference (OpenMP build) on `grid` with maxIters=iters; returns
    (compute seconds from its own report, GLUPS)."""
    if not os.path.exists(REF_OMP):
        raise RuntimeError(f"{REF_OMP} missing (built by oracle/build_ref.sh where /root/reference exists)")
    pfile, ofile = grid_files(grid)
    with open(pfile) as fh:
        lines = fh.read().split("\n")
    nx, ny = int(lines[0]), int(lines[1])
    lines[2] = str(iters)
    with tempfile.TemporaryDirectory() as tmp:
        pf = os.path.join(tmp, "sample.params")
        with open(pf, "w") as fh:
            fh.write("\n".join(lines))
        env = dict(os.environ, OMP_NUM_THREADS=str(threads), OMP_PROC_BIND="true", OMP_PLACES="cores")
        out = subprocess.run([REF_OMP, pf, ofile], cwd=tmp, env=env, check=True, capture_output=True, text=True).stdout
    secs = float(re.search(r"Elapsed Compute time:\s+([0-9.]+)", out).group(1))
    return secs, nx * ny * iters / secs / 1e9


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def reference_sample_iters(grid, threads, target_s):
    """Picks maxIters so that one sample takes about target_s seconds on this box."""
    probe = 100
    secs, _ = run_reference_sample(grid, probe, threads)
    return max(probe, int(probe * target_s / max(secs, 1e-3)))


def cpu_baseline(grid, target_s=12.0):
    threads = host_cores()
    iters = reference_sample_iters(grid, threads, target_s)
    secs, glups = run_reference_sample(grid, iters, threads)
    return {"value": glups, "unit": "GLUPS", "cores": threads, "kind": "reference",
            "sample": f"oracle/_ref/ref-omp (unmodified reference, gcc -O3 -fopenmp, OMP_PROC_BIND=true) on {grid} "
                      f"with maxIters={iters}: {secs:.2f} s compute"}


def reference_arm(args, rank, world):
    """bench.py --impl reference: the reference's own CPU implementation on this box's host cores."""
    if rank != 0:
        return 0
    grid = args.workload or "16384x1024"
    threads = host_cores()
    try:
        iters = reference_sample_iters(grid, threads, args.ref_seconds)
        for _ in range(args.warmup):
            run_reference_sample(grid, max(iters // 4, 50), threads)
        secs, updates = 0.0, 0.0
        for _ in range(args.steps):
            s, g = run_reference_sample(grid, iters, threads)
            secs += s
            updates += g * s
    except Exception as e:   # the oracle is built in-tree; report rather than crash the driver
        print(json.dumps({"impl": "reference", "unavailable": str(e)[:200]}))
        return 0
    glups = updates / secs
    nx, ny = (int(v) for v in grid.split("x"))
    line = {"impl": "reference", "metric": "GLUPS (lattice updates/sec)", "value": glups, "unit": "GLUPS",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": grid, "nx": nx, "ny": ny, "iters_per_step": iters,
                       "note": "bounded sample: maxIters reduced (LBM cost per timestep is constant)"},
            "cpu_baseline": {"value": glups, "unit": "GLUPS", "cores": threads, "kind": "reference",
                             "sample": f"oracle/_ref/ref-omp on {grid}, maxIters={iters} per step, {args.steps} steps"},
            "e2e": {"value": glups, "unit": "GLUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


# --------------------------------------------------------------------------- GPU arm
def flush_l2(torch, scratch):
    scratch.add_(1)          # 256 MiB read+write > 126 MB L2


def load_case(lbm, grid):
    """(params dict, int32 mask[ny, nx], obstacle-file line count) of a shipped grid (inputs/), or of a grid
    the reference does not ship (the MPI report's long-x / short-y 131072x128) by the wide-grid rule of
    tools/gen_inputs.py."""
    import numpy as np
    pfile, ofile = grid_files(grid)
    if os.path.exists(pfile):
        p = lbm.read_params(pfile)
        mask, lines = lbm.read_obstacles(ofile, p["nx"], p["ny"])
        return p, mask, lines
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import gen_inputs
    p = gen_inputs.synthetic_params(grid)
    cells = np.array(gen_inputs.obstacle_cells(grid), np.int64)
    mask = np.zeros((p["ny"], p["nx"]), np.int32)
    mask[cells[:, 1], cells[:, 0]] = 1
    return p, mask, int(cells.shape[0])


def make_lattice(lbm, torch, dist, case, cap, rank, world, **kw):
    """One lattice over `world` GPUs (one slab per rank, halo slots linked through CUDA IPC) or on this GPU."""
    ring = importlib.import_module("lattice-boltzmann_b200.ring")
    p, mask, lines = case
    if world > 1:
        lat = lbm.Lattice(p["nx"], p["ny"], cap, p["density"], p["accel"], p["omega"], mask, obstacle_lines=lines,
                          rank=rank, size=world, device=torch.cuda.current_device(), **kw)
        up, down, _ = ring.exchange_blobs(dist, torch, lat.rank_export(), rank, world, device="cuda")
        lat.rank_connect(up, down)
    else:
        lat = lbm.Lattice(p["nx"], p["ny"], cap, p["density"], p["accel"], p["omega"], mask, obstacle_lines=lines,
                          first_device=torch.cuda.current_device(), **kw)
    return lat


def close_lattice(lat, world, dist):
    """Teardown as include/lbm_b200.h prescribes for one process per GPU: disconnect, destroy - no barrier."""
    if world > 1:
        lat.rank_disconnect()
    lat.close()


def time_resident(lbm, torch, grid, iters, steps, warmup, rank, world, dist, scratch):
    """value leg: state resident in HBM, K steps of `iters` timesteps. Returns dict."""
    case = load_case(lbm, grid)
    p = case[0]
    lat = make_lattice(lbm, torch, dist, case, iters * (steps + warmup), rank, world)
    lat.sync()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    for _ in range(warmup):
        lat.run(iters)
        lat.sync()
    barrier()
    sampler = ClockSampler(torch.cuda.current_device())
    if rank == 0:
        sampler.start()
    dev_ms, launches = 0.0, 0
    t0 = time.perf_counter()
    for _ in range(steps):
        flush_l2(torch, scratch)
        barrier()
        lat.run(iters)
        lat.sync()
        ms = lat.last_run_ms()
        if world > 1:
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        dev_ms += ms
        launches += lat.last_run_launches()
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([launches], device="cuda", dtype=torch.int64)
        dist.all_reduce(t)
        launches = int(t.item())
    av = lat.av_vels(min(lat.steps_done, 8))
    desc = lat.describe()
    close_lattice(lat, world, dist)
    updates = float(p["nx"]) * p["ny"] * iters * steps
    return dict(glups=updates / (dev_ms * 1e-3) / 1e9, ms_per_step=dev_ms / steps, wall_s=wall, launches=launches,
                clocks=clocks, desc=desc, nx=p["nx"], ny=p["ny"], av0=float(av[0]) if len(av) else None)


def time_e2e(lbm, torch, grid, iters, steps, rank, world, dist):
    """e2e leg: host buffers in, host buffers out, through the public API, every step: lbm_create (obstacle
    int array from host memory -> bit mask upload, schedule upload) + lbm_run + lbm_get_av_vels + lbm_get_cells
    into the caller's (long-lived, pinned) result planes + destroy.  The result planes are allocated ONCE, as the
    reference allocates cells_global once in initialise() (d2q9-bgk.c:562-571)."""
    import numpy as np
    ring = importlib.import_module("lattice-boltzmann_b200.ring")
    case = load_case(lbm, grid)
    p, mask, lines = case
    rows = lbm.allocate_rows(p["ny"], world, rank)[1] if world > 1 else p["ny"]
    out = torch.empty((9, rows, p["nx"]), dtype=torch.float32, pin_memory=True).numpy()
    h2d = d2h = 0
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    ph = {"create": 0.0, "connect": 0.0, "run": 0.0, "download": 0.0, "reduce": 0.0, "destroy": 0.0}   # host seconds, this rank
    for _ in range(steps):
        ta = time.perf_counter()
        if world > 1:
            lat = lbm.Lattice(p["nx"], p["ny"], iters, p["density"], p["accel"], p["omega"], mask, obstacle_lines=lines,
                              rank=rank, size=world, device=torch.cuda.current_device())
            tb = time.perf_counter()
            up, down, _ = ring.exchange_blobs(dist, torch, lat.rank_export(), rank, world, device="cuda")
            lat.rank_connect(up, down)
        else:
            lat = lbm.Lattice(p["nx"], p["ny"], iters, p["density"], p["accel"], p["omega"], mask, obstacle_lines=lines,
                              first_device=torch.cuda.current_device())
            tb = time.perf_counter()
        tc = time.perf_counter()
        lat.run(iters)
        av = lat.av_vels()                   # waits for the run
        td = time.perf_counter()
        f = lat.cells(out=out)
        te = time.perf_counter()
        if world > 1:
            av = ring.reduce_av_vels(dist, torch, av, lat.inv_free_cells(), device="cuda")   # reference :445-446
        tf = time.perf_counter()
        a, b = lat.transfer_bytes()
        h2d += a
        d2h += b
        close_lattice(lat, world, dist)
        tg = time.perf_counter()
        for k, v in zip(ph, (tb - ta, tc - tb, td - tc, te - td, tf - te, tg - tf)):
            ph[k] += v / steps
        assert np.isfinite(av).all() and np.isfinite(f[:, ::64, ::64]).all()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    wall = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([wall, float(h2d), float(d2h)], device="cuda", dtype=torch.float64)
        tm = t.clone()
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        dist.all_reduce(t)
        wall, h2d, d2h = float(tm[0]), float(t[1]), float(t[2])
    updates = float(p["nx"]) * p["ny"] * iters * steps
    return dict(glups=updates / wall / 1e9, h2d=int(h2d / steps), d2h=int(d2h / steps),
                phases={k: round(v, 4) for k, v in ph.items()})


# --------------------------------------------------------------------------- parity evidence inside the bench
def plane_sha256(a):
    import hashlib
    import numpy as np
    return hashlib.sha256(np.ascontiguousarray(a, dtype="<f4").tobytes()).hexdigest()


def gather_planes(torch, dist, planes, case, rank, world):
    """Row slabs of 2-D planes -> the global planes on rank 0 (None elsewhere)."""
    import numpy as np
    lbm = importlib.import_module("lattice-boltzmann_b200")
    if world == 1:
        return planes
    ny = case[0]["ny"]
    counts = [lbm.allocate_rows(ny, world, r)[1] for r in range(world)]
    mx = max(counts)
    stack = np.zeros((len(planes), mx, planes[0].shape[1]), np.float32)
    for i, a in enumerate(planes):
        stack[i, :a.shape[0]] = a
    mine = torch.from_numpy(stack).cuda()
    got = [torch.empty_like(mine) for _ in range(world)] if rank == 0 else None
    dist.gather(mine, got, dst=0)
    if rank != 0:
        return None
    parts = [g.cpu().numpy() for g in got]
    return [np.concatenate([parts[r][i, :counts[r]] for r in range(world)], axis=0) for i in range(len(planes))]


def parity_case(lbm, torch, dist, grid, steps, rank, world):
    """One grid, `steps` timesteps on `world` GPUs:
      (1) [world > 1] every rank also runs the whole lattice on its own GPU and compares its slab's nine
          populations with those rows BIT FOR BIT (N-GPU == 1-GPU), av_vels after the one sum over ranks to 3e-7;
      (2) the four final-state planes are gathered on rank 0 and their SHA-256 compared with what the UNMODIFIED
          reference wrote for this input (tests/golden/ref_hash_<grid>_<steps>.json, a committed fixture made by
          tests/golden/make_golden.py from oracle/_ref/ref-serial)."""
    import numpy as np
    ring = importlib.import_module("lattice-boltzmann_b200.ring")
    case = load_case(lbm, grid)
    p = case[0]
    lat = make_lattice(lbm, torch, dist, case, steps, rank, world)
    lat.run(steps // 2)
    lat.run(steps - steps // 2)                     # a second launch re-arms tags and republishes the edge rows
    f = lat.cells()
    fs = lat.final_state()
    av = lat.av_vels()
    if world > 1:
        av = ring.reduce_av_vels(dist, torch, av, lat.inv_free_cells(), device="cuda")
    desc = lat.describe()
    row_start, rows = lat.row_start, lat.rows
    close_lattice(lat, world, dist)
    res = {"grid": grid, "steps": steps, "engine": desc}
    ok = True
    if world > 1:
        with lbm.Lattice(p["nx"], p["ny"], steps, p["density"], p["accel"], p["omega"], case[1], obstacle_lines=case[2],
                         first_device=torch.cuda.current_device()) as one:
            one.run(steps)
            f1 = one.cells()[:, row_start:row_start + rows]
            av1 = one.av_vels()
            res["single_gpu_engine"] = one.describe()
        bad = int(np.count_nonzero(f.view(np.uint32) != f1.view(np.uint32)))
        av_err = float(np.max(np.abs(av - av1) / np.abs(av1)))
        t = torch.tensor([bad, int(av_err > 3e-7)], device="cuda", dtype=torch.int64)
        dist.all_reduce(t)
        res["n_gpu_equals_1_gpu"] = int(t[0].item()) == 0
        res["mismatching_values"] = int(t[0].item())
        res["av_vels_max_rel_diff"] = av_err
        ok = ok and res["n_gpu_equals_1_gpu"] and int(t[1].item()) == 0
    planes = gather_planes(torch, dist, list(fs), case, rank, world)
    gpath = os.path.join(ROOT, "tests", "golden", f"ref_hash_{grid}_{steps}.json")
    if rank == 0 and os.path.exists(gpath):
        with open(gpath) as fh:
            g = json.load(fh)
        same = all(plane_sha256(a) == g["sha256_" + n] for n, a in zip(("ux", "uy", "u", "pressure"), planes))
        ref_av = np.array([float(v) for v in g["av_vels"]], np.float32)
        res["final_state_sha256_equals_reference"] = bool(same)
        res["av_vels_vs_reference_max_rel_diff"] = float(np.max(np.abs(av - ref_av) / np.abs(ref_av)))
        # the reference sums |u| sequentially in fp32 over up to 16.8 M cells (d2q9-bgk.c:335): its own rounding error
        # reaches 5e-4 there (tests/test_gpu_parity.py holds the device's double-precision sum to 3e-7 of the exact sum)
        ok = ok and same and res["av_vels_vs_reference_max_rel_diff"] <= 1e-3
    if world > 1:
        t = torch.tensor([int(ok)], device="cuda", dtype=torch.int64)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        ok = bool(t.item())
    res["ok"] = ok
    return res


def parity_check(lbm, torch, dist, grid, rank, world):
    """Correctness record printed with every bench line (the GPU test box has one GPU, so at N > 1 this is the
    driver-visible evidence that the row-slab path computes what one GPU and the unmodified reference compute):
    the bench workload for 20 steps against the reference's hashes, for 200 steps N-GPU against 1-GPU, and the
    periodic 128x256 grid (the ring-wrap link between the last and the first slab) for 200 steps against both."""
    cases = [parity_case(lbm, torch, dist, grid, 20, rank, world)]
    if world > 1:
        cases.append(parity_case(lbm, torch, dist, grid, 200, rank, world))
    cases.append(parity_case(lbm, torch, dist, "128x256", 200, rank, world))
    out = {"ok": all(c["ok"] for c in cases), "cases": cases}
    if world > 1:
        out["n_gpu_equals_1_gpu"] = all(c.get("n_gpu_equals_1_gpu", True) for c in cases)
        out["steps"] = 200
        out["grid"] = grid
    return out


def also_leg(lbm, torch, dist, scratch, grid, rank, world, iters=None, steps=3, note="", vs_single=False, e2e=True):
    """One more configuration under "also": device-timed value, roofline fraction per GPU, optional e2e, and
    (vs_single, N > 1) the same lattice on ONE GPU in the same run, which gives the N-GPU efficiency."""
    hi = iters or load_case(lbm, grid)[0]["max_iters"]
    r = time_resident(lbm, torch, grid, hi, steps, 3, rank, world, dist, scratch)
    leg = {"value": r["glups"], "unit": "GLUPS", "iters_per_step": hi, "steps": steps, "ms_per_step": r["ms_per_step"],
           "us_per_timestep": r["ms_per_step"] * 1e3 / hi,
           "roofline_frac": r["glups"] * BYTES_PER_UPDATE / world / hbm_peak()[0], "gpu_launches": r["launches"],
           "engine": r["desc"], "note": note}
    if e2e:
        e = time_e2e(lbm, torch, grid, hi, 3, rank, world, dist)
        leg["e2e"] = e["glups"]
        leg["e2e_phases_s"] = e["phases"]
    if vs_single and world > 1:
        if rank == 0:
            one = time_resident(lbm, torch, grid, hi, steps, 3, 0, 1, None, scratch)
            leg["single_gpu_value"] = one["glups"]
            leg["single_gpu_engine"] = one["desc"]
            leg["efficiency_vs_single_gpu"] = r["glups"] / (world * one["glups"])
        dist.barrier()
    return leg


def cli_leg(grid):
    """Wall clock of the drop-in program itself, `./d2q9-bgk <params> <obstacles>` (d2q9-bgk.c:166-232), on one GPU:
    process start to exit, split by its own report (init / compute / collate) and LBM_REPORT's output-file times."""
    exe = os.path.join(ROOT, "d2q9-bgk")
    pfile, ofile = grid_files(grid)
    with tempfile.TemporaryDirectory() as tmp:
        t0 = time.perf_counter()
        r = subprocess.run([exe, pfile, ofile], cwd=tmp, capture_output=True, text=True, env=dict(os.environ, LBM_REPORT="1"))
        wall = time.perf_counter() - t0
        if r.returncode != 0:
            return {"error": r.stderr[-200:]}
        size = os.path.getsize(os.path.join(tmp, "final_state.dat"))
    leg = {"grid": grid, "process_wall_s": round(wall, 3), "final_state_bytes": size}
    for name in ("Init", "Compute", "Collate", "Total"):
        m = re.search(rf"Elapsed {name} time:\s+([0-9.]+)", r.stdout)
        leg[name.lower() + "_s"] = float(m.group(1)) if m else None
    m = re.search(r"output: final_state.dat ([0-9.]+) s \(format ([0-9.]+) s.*?\), av_vels.dat ([0-9.]+) s", r.stderr)
    if m:
        leg["write_final_state_s"], leg["format_s"], leg["write_av_vels_s"] = (float(v) for v in m.groups())
    return leg


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, help="grid name, e.g. 1024x1024 (default: by --gpus)")
    ap.add_argument("--iters", type=int, default=0, help="timesteps per step (default: the workload's maxIters)")
    ap.add_argument("--ref-seconds", type=float, default=12.0, help="CPU seconds per reference sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-also", action="store_true", help="skip the extra configurations under \"also\"")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity_check record")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return reference_arm(args, rank, world)
    if args.warmup < 3:
        args.warmup = 3

    # stdout carries exactly ONE line, the JSON record: file descriptor 1 is pointed at stderr for the rest of the run
    # (NCCL prints its version banner and NCCL_DEBUG output on stdout from C; NCCL_DEBUG itself is left alone so that
    # whoever launched the bench can read NCCL's log), and the record is written to the saved descriptor at the end.
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    import torch
    lbm = importlib.import_module("lattice-boltzmann_b200")
    if not torch.cuda.is_available() or lbm.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    dist = None
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    if world != args.gpus:
        raise SystemExit(f"bench.py: --gpus {args.gpus} needs {args.gpus} ranks (torchrun), got WORLD_SIZE={world}")
    scratch = torch.zeros(64 * 1024 * 1024, dtype=torch.float32, device="cuda")

    grid = args.workload or "16384x1024"
    iters = args.iters or load_case(lbm, grid)[0]["max_iters"]
    parity = None if args.no_parity else parity_check(lbm, torch, dist, grid, rank, world)
    if parity is not None and not parity["ok"]:
        if rank == 0:
            os.write(json_fd, (json.dumps({"error": "parity_check failed", "parity_check": parity}) + "\n").encode())
        return 1
    res = time_resident(lbm, torch, grid, iters, args.steps, args.warmup, rank, world, dist, scratch)
    e2e = time_e2e(lbm, torch, grid, iters, max(1, min(args.steps, 3)), rank, world, dist)
    also = {}
    if not args.workload and not args.no_also:
        # BASELINE.json's other measured configurations: the single-GPU roofline headline at N = 1 (plus 2048x1024,
        # whose A/B set is just above L2, and the drop-in program's own wall clock), the weak-scaling pair (2.1 M cells
        # per GPU, as 16384x1024 at 8) at N = 2 and N = 4, and the few-rows-per-GPU regime at N = 8
        if world == 1:
            also["1024x1024"] = also_leg(lbm, torch, dist, scratch, "1024x1024", rank, world, steps=max(3, args.steps),
                                         note="BASELINE.json's single-GPU roofline headline; 75 MB A/B set is L2-resident inside a step, so the "
                                              "algorithmic rate may exceed the HBM roofline (ncu: 21 B of DRAM traffic per update)")
            also["2048x1024"] = also_leg(lbm, torch, dist, scratch, "2048x1024", rank, world, iters=5000, e2e=False,
                                         note="A/B working set 151 MB, just above the 126 MB L2 (the case an in-place AA scheme would target)")
            for g in ("128x128", "256x256"):
                also[g] = also_leg(lbm, torch, dist, scratch, g, rank, world, e2e=False,
                                   note="BASELINE.json check grid; small-lattice engine, latency-bound (L2-resident)")
            try:
                also["cli"] = cli_leg(grid)
            except Exception as e:      # e.g. the C driver was not built: the headline line must still be printed
                also["cli"] = {"error": str(e)[:200]}
        elif world in (2, 4):
            head = {2: "4096x1024", 4: "8192x1024"}[world]
            also[head] = also_leg(lbm, torch, dist, scratch, head, rank, world,
                                  note="BASELINE.json's weak-scaling configuration for this GPU count (2.1 M cells per GPU)")
        elif world == 8:
            also["131072x128"] = also_leg(lbm, torch, dist, scratch, "131072x128", rank, world, iters=2000, vs_single=True, e2e=False,
                                          note="the reference MPI report's long-x / short-y case (report_mpi/report.tex:837-859): 16 rows per GPU")
            also["1024x1024"] = also_leg(lbm, torch, dist, scratch, "1024x1024", rank, world, vs_single=True, e2e=False,
                                         note="small lattice on many GPUs: 128 rows x 1024 cells per GPU, latency-bound")
    if rank == 0:
        peak, how = hbm_peak()
        achieved = res["glups"] * BYTES_PER_UPDATE / world      # GB/s per GPU
        line = {"metric": "GLUPS (lattice updates/sec)", "value": res["glups"], "unit": "GLUPS", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": res["ms_per_step"], "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": grid, "nx": res["nx"], "ny": res["ny"], "iters_per_step": iters,
                           "parallelism": f"row-slabs x{world}" if world > 1 else "single GPU",
                           "engine": res["desc"],
                           "l2": "working set 1.2 GB >> 126 MB L2 (inputs larger than L2); additionally a 256 MiB flush between steps",
                           "inputs": "shipped-format params/obstacles regenerated by tools/gen_inputs.py (no RNG)"},
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": measured_traffic(grid, float(res["nx"]) * res["ny"] * iters / world),
                             "algorithmic_bytes_per_launch": BYTES_PER_UPDATE * res["nx"] * res["ny"] * iters / world,
                             "peak_source": how,
                             "note": "achieved = 72 B x cell updates of one persistent launch / its CUDA-event time, per GPU; "
                                     "traffic = ncu dram bytes per cell update (profiles/traffic.json) x updates per launch per GPU"},
                "e2e": {"value": e2e["glups"], "unit": "GLUPS", "h2d_bytes_per_step": e2e["h2d"], "d2h_bytes_per_step": e2e["d2h"],
                        "phases_s": e2e["phases"],
                        "note": "lbm_create + lbm_run + lbm_get_av_vels + lbm_get_cells + lbm_destroy per step, host buffers "
                                "(obstacle ints in pageable memory, result planes in the caller's long-lived pinned buffer)"},
                "gpu_launches": res["launches"], "clocks": res["clocks"], "wall_s_timed_region": res["wall_s"]}
        if parity is not None:
            line["parity_check"] = parity
        if also:
            line["also"] = also
        if world == 1 and not args.no_cpu_baseline:
            try:
                line["cpu_baseline"] = cpu_baseline(grid, args.ref_seconds)
            except Exception as e:
                line["cpu_baseline"] = {"value": None, "unit": "GLUPS", "cores": host_cores(), "kind": "reference",
                                        "sample": f"failed: {str(e)[:160]}"}
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())

#!/usr/bin/env python3
"""Generates the fixtures under tests/golden/ (run in the BUILD container only).

Needs /root/reference (read-only) and oracle/_ref/ref-serial (oracle/build_ref.sh).
Nothing here runs on the GPU box; the .npz files it writes are committed.

  shipped_<g>.npz        the reference's own known answers check/<g>.{av_vels,final_state}.dat
                         (the columns check/check.py:60-61 reads: av_vels col 1, final_state
                         cols 0,1,5), stored as float64 exactly as np.loadtxt parses them (the shipped
                         values are not fp32-representable: they came from a different build).
  ref_debug_<g>_<n>.txt  the per-step lines printed by the reference built with -DDEBUG (ref-debug)
  ref_full_<g>.npz       outputs of the UNMODIFIED reference (ref-serial, gcc -O3) on the shipped
                         inputs at full length: av_vels, u_x, u_y, |u|, pressure, Reynolds line.
                         Fills the two final_state goldens missing from the reference mount
                         (256x256, 1024x1024).  1024x1024 keeps pressure and |u| only (size).
  ref_short_<g>_<n>.npz  the same binary with maxIters cut to n: tight-tolerance target for the
                         CUDA path (fp32 rounding differences have no time to amplify).

  ref_hash_<g>_<n>.json  BASELINE.json's large configurations (4096x1024, 8192x1024, 16384x1024 - the bench
                         headline) and the periodic 128x256 grid, maxIters cut to n: SHA-256 of the fp32
                         u_x / u_y / |u| / pressure planes the unmodified reference wrote, plus its av_vels
                         values.  The planes are hundreds of MB, the fixture a few hundred bytes; the GPU
                         test hashes lbm_get_final_state's output (bit-exact or nothing), bench.py's
                         parity_check does the same at N > 1.  (--hashes-only makes just these.)

Full runs take 15 s .. 15 min each; pass --cache DIR to reuse earlier outputs
(DIR/<g>/{av_vels.dat,final_state.dat,stdout.txt}).
"""
import argparse
import os
import re
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "ref-serial")

SHIPPED = {"128x128": True, "128x256": True, "256x256": False, "1024x1024": False}  # has final_state?
SHORT = {"128x128": 1000, "128x256": 1000, "256x256": 1000, "1024x1024": 200, "2048x1024": 100}


DEBUG_CASE = ("128x256", 25)
HASHED = {"4096x1024": 20, "8192x1024": 20, "16384x1024": 20, "128x256": 200}


def load_av(path):
    return np.loadtxt(path, usecols=[1])


def f32_exact(a64, what):
    a32 = a64.astype(np.float32)
    bad = sum(("%.12E" % v) != ("%.12E" % w) for v, w in zip(a64.ravel()[:5000], a32.ravel()[:5000]))
    if bad:
        raise SystemExit(f"{what}: float32 does not reproduce the printed values ({bad} differ)")
    return a32


def shipped():
    for g, has_fs in SHIPPED.items():
        out = {"av_vels": load_av(f"{REF}/check/{g}.av_vels.dat")}
        if has_fs:
            fs = np.loadtxt(f"{REF}/check/{g}.final_state.dat", usecols=[0, 1, 5])
            nx, ny = (int(s) for s in g.split("x"))
            assert fs.shape[0] == nx * ny
            assert np.array_equal(fs[:, 0], np.tile(np.arange(nx), ny))
            assert np.array_equal(fs[:, 1], np.repeat(np.arange(ny), nx))
            out["pressure"] = fs[:, 2].reshape(ny, nx)
        np.savez_compressed(os.path.join(HERE, f"shipped_{g}.npz"), **out)
        print("wrote shipped", g, {k: v.shape for k, v in out.items()})


def run_ref(g, iters, workdir, binary=None):
    """Runs ref-serial (or `binary`) on grid g (maxIters overridden if iters) inside workdir."""
    with open(os.path.join(ROOT, "inputs", f"input_{g}.params")) as fh:
        lines = fh.read().split("\n")
    if iters:
        lines[2] = str(iters)
    pfile = os.path.join(workdir, "in.params")
    with open(pfile, "w") as fh:
        fh.write("\n".join(lines))
    res = subprocess.run([binary or REF_BIN, pfile, os.path.join(ROOT, "inputs", f"obstacles_{g}.dat")],
                         cwd=workdir, check=True, capture_output=True, text=True)
    with open(os.path.join(workdir, "stdout.txt"), "w") as fh:
        fh.write(res.stdout)


def collect(g, d, fields):
    nx, ny = (int(s) for s in g.split("x"))
    out = {"av_vels": f32_exact(load_av(os.path.join(d, "av_vels.dat")), g)}
    cols = {"ux": 2, "uy": 3, "u": 4, "pressure": 5}
    fs = np.loadtxt(os.path.join(d, "final_state.dat"), usecols=[cols[f] for f in fields] + [6])
    for i, f in enumerate(fields):
        out[f] = f32_exact(fs[:, i], f"{g}.{f}").reshape(ny, nx)
    out["obstacles"] = fs[:, -1].astype(np.uint8).reshape(ny, nx)
    with open(os.path.join(d, "stdout.txt")) as fh:
        m = re.search(r"Reynolds number:\s+(\S+)", fh.read())
    out["reynolds"] = np.array(m.group(1))
    return out


def plane_sha256(a):
    import hashlib
    return hashlib.sha256(np.ascontiguousarray(a, dtype="<f4").tobytes()).hexdigest()


def hashes():
    """ref_hash_<g>_<n>.json: see the module docstring.  final_state.dat of 16384x1024 is 1.4 GB of text:
    parsed with pandas (float64, exact), then checked to be fp32-representable before hashing."""
    import json
    import pandas as pd
    for g, n in HASHED.items():
        nx, ny = (int(s) for s in g.split("x"))
        with tempfile.TemporaryDirectory() as tmp:
            run_ref(g, n, tmp)
            fs = pd.read_csv(os.path.join(tmp, "final_state.dat"), sep=" ", header=None,
                             names=["x", "y", "ux", "uy", "u", "pressure", "obst"], float_precision="round_trip")
            av = f32_exact(load_av(os.path.join(tmp, "av_vels.dat")), g)
            with open(os.path.join(tmp, "stdout.txt")) as fh:
                re_line = re.search(r"Reynolds number:\s+(\S+)", fh.read()).group(1)
        assert len(fs) == nx * ny and int(fs["x"].iloc[-1]) == nx - 1 and int(fs["y"].iloc[-1]) == ny - 1
        out = {"grid": g, "steps": n, "made_by": "oracle/_ref/ref-serial (unmodified reference, gcc -O3)",
               "dtype": "<f4, row-major [ny][nx]", "reynolds": re_line,
               "av_vels": ["%.9e" % v for v in av.tolist()], "blocked_cells": int(fs["obst"].sum())}
        for f in ("ux", "uy", "u", "pressure"):
            a64 = fs[f].to_numpy()
            a32 = f32_exact(a64, f"{g}.{f}")            # %.12E of the fp32 value == the printed text (sampled)
            # ... and for every value: 13 printed digits pin the fp32 value (9 suffice), so the nearest
            # fp32 must lie within the print rounding (half a unit of the 13th digit) of the text
            if not np.all(np.abs(a32.astype(np.float64) - a64) <= 1e-12 * np.abs(a64)):
                raise SystemExit(f"{g}.{f}: printed values are not fp32 values")
            out["sha256_" + f] = plane_sha256(a32.reshape(ny, nx))
        with open(os.path.join(HERE, f"ref_hash_{g}_{n}.json"), "w") as fh:
            json.dump(out, fh, indent=1)
        print("wrote ref_hash", g, n, out["sha256_pressure"][:16], flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--hashes-only", action="store_true", help="only the ref_hash_* fixtures of the large grids")
    ap.add_argument("--cache", help="directory with earlier full-length ref-serial outputs")
    ap.add_argument("--skip-full", action="store_true")
    args = ap.parse_args()
    if not (os.path.isdir(REF) and os.path.exists(REF_BIN)):
        raise SystemExit("needs /root/reference and oracle/_ref/ref-serial (run oracle/build_ref.sh)")
    if args.hashes_only:
        hashes()
        return 0
    shipped()
    if not args.skip_full:
        for g in SHIPPED:
            fields = ["u", "pressure"] if g == "1024x1024" else ["ux", "uy", "u", "pressure"]
            d = os.path.join(args.cache, g) if args.cache else None
            if d and os.path.exists(os.path.join(d, "stdout.txt")):
                out = collect(g, d, fields)
            else:
                with tempfile.TemporaryDirectory() as tmp:
                    run_ref(g, None, tmp)
                    out = collect(g, tmp, fields)
            np.savez_compressed(os.path.join(HERE, f"ref_full_{g}.npz"), **out)
            print("wrote ref_full", g, str(out["reynolds"]))
    for g, n in SHORT.items():
        fields = ["u", "pressure"] if int(g.split("x")[0]) >= 1024 else ["ux", "uy", "u", "pressure"]
        with tempfile.TemporaryDirectory() as tmp:
            run_ref(g, n, tmp)
            out = collect(g, tmp, fields)
        np.savez_compressed(os.path.join(HERE, f"ref_short_{g}_{n}.npz"), **out)
        print("wrote ref_short", g, n)
    # the reference's -DDEBUG build (d2q9-bgk.c:198-202): per-step "av velocity" / "tot density" lines
    g, n = DEBUG_CASE
    with tempfile.TemporaryDirectory() as tmp:
        run_ref(g, n, tmp, os.path.join(ROOT, "oracle", "_ref", "ref-debug"))
        with open(os.path.join(tmp, "stdout.txt")) as fh:
            dbg = [l for l in fh.read().splitlines(True) if re.match(r"==timestep|av velocity|tot density", l)]
    with open(os.path.join(HERE, f"ref_debug_{g}_{n}.txt"), "w") as fh:
        fh.writelines(dbg)
    print("wrote ref_debug", g, n, len(dbg), "lines")
    hashes()
    return 0


if __name__ == "__main__":
    sys.exit(main())

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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cache", help="directory with earlier full-length ref-serial outputs")
    ap.add_argument("--skip-full", action="store_true")
    args = ap.parse_args()
    if not (os.path.isdir(REF) and os.path.exists(REF_BIN)):
        raise SystemExit("needs /root/reference and oracle/_ref/ref-serial (run oracle/build_ref.sh)")
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
    return 0


if __name__ == "__main__":
    sys.exit(main())

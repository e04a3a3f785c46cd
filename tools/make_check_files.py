#!/usr/bin/env python3
"""Writes check/<grid>.av_vels.dat and check/<grid>.final_state.dat - the known answers `make check`
compares against (reference Makefile:13-14, 23-24) - from the committed fixtures tests/golden/shipped_<grid>.npz
(the reference's own check/*.dat values as parsed by np.loadtxt, see tests/golden/make_golden.py).  Only the
columns check/check.py:60-61 reads are meaningful: av_vels col 1; final_state cols 0, 1 and 5 (pressure); the
velocity columns are written as zeros and the obstacle column from the obstacle input file."""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("grids", nargs="*", default=["128x128", "128x256", "256x256", "1024x1024"])
    ap.add_argument("--out", default=os.path.join(ROOT, "check"))
    args = ap.parse_args()
    os.makedirs(args.out, exist_ok=True)
    for g in args.grids:
        z = np.load(os.path.join(ROOT, "tests", "golden", f"shipped_{g}.npz"))
        with open(os.path.join(args.out, f"{g}.av_vels.dat"), "w") as fh:
            fh.write("".join("%d:\t%.12E\n" % (i, v) for i, v in enumerate(z["av_vels"])))
        if "pressure" not in z:
            continue                                   # the reference ships no final_state for this grid
        pr = z["pressure"]
        ny, nx = pr.shape
        obst = np.zeros((ny, nx), np.int64)
        with open(os.path.join(ROOT, "inputs", f"obstacles_{g}.dat")) as fh:
            for line in fh:
                x, y, _ = line.split()
                obst[int(y), int(x)] = 1
        with open(os.path.join(args.out, f"{g}.final_state.dat"), "w") as fh:
            for jj in range(ny):
                fh.write("".join("%d %d %.12E %.12E %.12E %.12E %d\n" % (ii, jj, 0.0, 0.0, 0.0, pr[jj, ii], obst[jj, ii])
                                 for ii in range(nx)))
        print("wrote", g)
    return 0


if __name__ == "__main__":
    sys.exit(main())

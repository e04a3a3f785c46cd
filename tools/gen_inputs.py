#!/usr/bin/env python3
"""Regenerates the benchmark inputs under inputs/ from their generating rules.

The reference ships `input_<g>.params` (7 lines: nx ny maxIters reynolds_dim density accel
omega; read at d2q9-bgk.c:522-528) and `obstacles_<g>.dat` (lines "x y 1"; read at
d2q9-bgk.c:628-641) for eight grids.  /root/reference does not exist on the GPU box, so the
same inputs are produced here from the rule each file follows (box walls as runs of cells,
optional full-height interior walls; cf. generate_obstacles.py:6-19).  Line ORDER and
DUPLICATES (corner cells appear twice) are kept because the reference's av_vels denominator
counts obstacle-file LINES (d2q9-bgk.c:626,640,642).

`--verify REFDIR` checks byte equality with the reference's shipped files.
"""
import argparse
import os
import sys

# grid -> (nx, ny, maxIters, reynolds_dim, density, accel, omega)
PARAMS = {
    "128x128":    (128, 128, 40000, 10, "0.1", "0.005", "1.85"),
    "128x256":    (128, 256, 40000, 10, "0.1", "0.005", "1.85"),
    "256x256":    (256, 256, 80000, 10, "0.1", "0.005", "1.85"),
    "1024x1024":  (1024, 1024, 20000, 10, "0.1", "0.01", "1.85"),
    "2048x1024":  (2048, 1024, 20000, 10, "0.1", "0.01", "1.85"),
    "4096x1024":  (4096, 1024, 20000, 10, "0.1", "0.01", "1.85"),
    "8192x1024":  (8192, 1024, 20000, 10, "0.1", "0.01", "1.85"),
    "16384x1024": (16384, 1024, 20000, 10, "0.1", "0.01", "1.85"),
}


def hrun(y, nx):
    return [(x, y) for x in range(nx)]


def vrun(x, ny):
    return [(x, y) for y in range(ny)]


def obstacle_cells(grid):
    """Ordered list of (x, y) lines of obstacles_<grid>.dat.  Grids the reference does not ship (e.g. its MPI
    report's long-x / short-y case 131072x128, report_mpi/report.tex:837-859) follow the wide-grid rule."""
    nx, ny = PARAMS[grid][:2] if grid in PARAMS else (int(v) for v in grid.split("x"))
    if grid == "128x256":          # open (periodic) top/bottom, a wall across y=127, side walls
        return hrun(127, nx) + vrun(0, ny) + vrun(nx - 1, ny)
    if nx <= 1024:                 # closed box: bottom, top, left, right
        cells = hrun(0, nx) + hrun(ny - 1, nx) + vrun(0, ny) + vrun(nx - 1, ny)
        if grid == "1024x1024":
            cells += vrun(341, ny)  # one interior wall
        return cells
    # wide grids: bottom, top, right, left, then walls at nx/4 and nx/2
    return (hrun(0, nx) + hrun(ny - 1, nx) + vrun(nx - 1, ny) + vrun(0, ny)
            + vrun(nx // 4, ny) + vrun((nx // 4) * 2, ny))


def synthetic_params(grid):
    """Params of a grid outside the shipped set: the shipped wide grids' values."""
    nx, ny = (int(v) for v in grid.split("x"))
    return dict(nx=nx, ny=ny, max_iters=20000, reynolds_dim=10, density=0.1, accel=0.01, omega=1.85)


def params_text(grid):
    return "".join(f"{v}\n" for v in PARAMS[grid])


def obstacles_text(grid):
    return "".join(f"{x} {y} 1\n" for x, y in obstacle_cells(grid))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "inputs"))
    ap.add_argument("--verify", metavar="REFDIR", help="compare with the reference's shipped inputs")
    args = ap.parse_args()
    os.makedirs(args.out, exist_ok=True)
    bad = 0
    for g in PARAMS:
        for name, text in ((f"input_{g}.params", params_text(g)), (f"obstacles_{g}.dat", obstacles_text(g))):
            with open(os.path.join(args.out, name), "w") as fh:
                fh.write(text)
            if args.verify:
                with open(os.path.join(args.verify, name)) as fh:
                    same = fh.read() == text
                print(f"{name}: {'identical' if same else 'DIFFERS'}")
                bad += not same
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())

#!/usr/bin/env python3
"""Tolerance checker for av_vels.dat / final_state.dat with the CLI and the pass/fail rule of
the reference's check/check.py (which is used UNCHANGED wherever the reference tree exists;
this restatement serves the GPU box, where it does not).

Rule (check/check.py:57-88,136-151): load av_vels column 1 and final_state columns 0,1,5;
coordinates must be identical and the step counts equal; with diff = ref - sim the percentage
is 100*diff/(ref - diff); a file fails if the largest |percentage| is non-finite or exceeds
--tolerance (default 1 %).  Exit status 0 = both passed, 1 otherwise.
"""
import argparse
import sys

import numpy as np


def compare(ref, sim):
    diff = ref - sim
    with np.errstate(divide="ignore", invalid="ignore"):
        pcnt = 100.0 * (diff / (ref - diff))
    worst = int(np.argmax(np.abs(pcnt)))
    return dict(worst=worst, diff=diff[worst], pcnt=pcnt[worst], sim=sim[worst], ref=ref[worst],
                total=float(np.sum(np.abs(diff))))


def main():
    ap = argparse.ArgumentParser(description="LBM result checker", fromfile_prefix_chars="@")
    ap.add_argument("--tolerance", nargs=1, default=[1], type=float)
    ap.add_argument("--ref-av-vels-file", nargs=1, required=True)
    ap.add_argument("--ref-final-state-file", nargs=1, required=True)
    ap.add_argument("--av-vels-file", nargs=1, required=True)
    ap.add_argument("--final-state-file", nargs=1, required=True)
    a = ap.parse_args()
    tol = a.tolerance[0]

    av_ref = np.loadtxt(a.ref_av_vels_file[0], usecols=[1])
    av_sim = np.loadtxt(a.av_vels_file[0], usecols=[1])
    fs_ref = np.loadtxt(a.ref_final_state_file[0], usecols=[0, 1, 5])
    fs_sim = np.loadtxt(a.final_state_file[0], usecols=[0, 1, 5])

    if fs_ref.shape != fs_sim.shape or np.any(fs_ref[:, :2] != fs_sim[:, :2]):
        print("Final state files coordinates were not the same")
        return 1
    if av_ref.size != av_sim.size:
        print("Different number of steps in av_vels files")
        return 1

    av = compare(av_ref, av_sim)
    print(f"Total difference in av_vels : {av['total']:.12E}")
    print(f"Biggest difference (at step {av['worst']:d}) : {av['diff']:.12E}")
    print(f"  {av['sim']:.12E} vs. {av['ref']:.12E} = {av['pcnt']:.2g}%\n")
    fs = compare(fs_ref[:, 2], fs_sim[:, 2])
    jj, ii = int(fs_sim[fs["worst"], 0]), int(fs_sim[fs["worst"], 1])
    print(f"Total difference in final_state : {fs['total']:.12E}")
    print(f"Biggest difference (at coord ({jj:d},{ii:d})) : {fs['diff']:.12E}")
    print(f"  {fs['sim']:.12E} vs. {fs['ref']:.12E} = {fs['pcnt']:.2g}%\n")

    fs_failed = (not np.isfinite(fs["pcnt"])) or abs(fs["pcnt"]) > tol
    av_failed = (not np.isfinite(av["pcnt"])) or abs(av["pcnt"]) > tol
    if fs_failed:
        print("final state failed check")
    if av_failed:
        print("av_vels failed check")
    if fs_failed or av_failed:
        return 1
    print("Both tests passed!")
    return 0


if __name__ == "__main__":
    sys.exit(main())

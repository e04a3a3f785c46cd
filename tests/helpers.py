"""Shared helpers for the parity tests (test infrastructure)."""
import hashlib
import json
import os

import numpy as np

import oracle_lib

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def plane_sha256(a):
    """SHA-256 of a plane as little-endian fp32, row-major (how tests/golden/make_golden.py hashes
    what the unmodified reference wrote)."""
    return hashlib.sha256(np.ascontiguousarray(a, dtype="<f4").tobytes()).hexdigest()


def load_ref_hash(grid, steps):
    """tests/golden/ref_hash_<grid>_<steps>.json: hashes of the reference's final-state planes + its av_vels."""
    with open(os.path.join(GOLDEN_DIR, f"ref_hash_{grid}_{steps}.json")) as fh:
        g = json.load(fh)
    g["av_vels"] = np.array([float(v) for v in g["av_vels"]], np.float32)
    return g


def assert_final_state_matches_ref_hash(g, ux, uy, u, pr):
    for name, a in (("ux", ux), ("uy", uy), ("u", u), ("pressure", pr)):
        assert plane_sha256(a) == g["sha256_" + name], f"{g['grid']}: {name} plane differs from the reference's"


def load_grid(grid):
    pfile, ofile = oracle_lib.grid_files(grid)
    p = oracle_lib.read_params(pfile)
    mask, lines = oracle_lib.read_obstacles(ofile, p["nx"], p["ny"])
    return p, mask, lines


def random_case(nx, ny, seed, walls=True, blocks=6):
    """Seeded obstacle mask + perturbed equilibrium-like positive lattice."""
    rng = np.random.default_rng(seed)
    mask = np.zeros((ny, nx), np.int32)
    if walls:
        mask[0, :] = mask[-1, :] = 1
        mask[:, 0] = mask[:, -1] = 1
    for _ in range(blocks):
        w, h = rng.integers(1, max(2, nx // 8)), rng.integers(1, max(2, ny // 8))
        x, y = rng.integers(0, nx - w), rng.integers(0, ny - h)
        mask[y:y + h, x:x + w] = 1
    mask[ny - 2, nx // 2: nx // 2 + 3] = 0   # keep some fluid on the accelerated row
    w = np.array([4 / 9] + [1 / 9] * 4 + [1 / 36] * 4, np.float32) * np.float32(0.1)
    f = (w[:, None, None] * (1.0 + 0.2 * rng.standard_normal((9, ny, nx)))).astype(np.float32)
    return mask, np.abs(f) + np.float32(1e-4)


def check_py_max_pcnt(ref, sim):
    """check/check.py:83-88: diff = ref - sim; pcnt = 100*diff/(ref - diff); max |pcnt|."""
    ref = np.asarray(ref, np.float64).ravel()
    sim = np.asarray(sim, np.float64).ravel()
    diff = ref - sim
    with np.errstate(divide="ignore", invalid="ignore"):
        pc = 100.0 * (diff / (ref - diff))
    return float(np.max(np.abs(pc)))

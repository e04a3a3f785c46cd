"""-m gpu test of lbm_diagnostics_noalloc, the entry point that allocates nothing (its scratch is a fixed file-scope
device array; include/lbm_b200.h).  In a file of its own that pytest collects last.

Through the C ABI, on two calls in a row (the scratch array keeps the previous call's contents) and on a second lattice
of another shape: the results must equal the allocating entry points' (lbm_av_velocity, lbm_total_density) to 3e-7
relative - the tolerance the other tests use for double-precision sums that are rounded to float (both paths add the same
per-cell fp32 values in double, in different orders).  A slab with more rows than the scratch holds must be rejected."""
import numpy as np
import pytest

from helpers import load_grid, random_case

pytestmark = pytest.mark.gpu

RTOL = 3e-7


def _both(lat):
    return (np.float32(lat.av_velocity()), np.float32(lat.total_density())), lat.diagnostics_noalloc()


def test_noalloc_diagnostics_equal_the_allocating_entry_points_on_two_calls_in_a_row(lbm):
    p, mask, lines = load_grid("128x256")
    with lbm.Lattice(p["nx"], p["ny"], 300, p["density"], p["accel"], p["omega"], mask, obstacle_lines=lines) as lat:
        lat.run(100)
        want1, got1 = _both(lat)
        lat.run(200)                              # the state changes between the two calls
        want2, got2 = _both(lat)
        again = lat.diagnostics_noalloc()         # and a third call on an unchanged state: deterministic
    np.testing.assert_allclose(got1, want1, rtol=RTOL)
    np.testing.assert_allclose(got2, want2, rtol=RTOL)
    assert want1[0] != want2[0] and got1[0] != got2[0]          # the second call did not return the first call's result
    assert again == got2
    # a lattice of another shape (fewer rows: stale entries of the larger one stay in the shared array) and a random state
    mask2, f0 = random_case(512, 40, seed=31, walls=True)
    with lbm.Lattice(512, 40, 10, 0.1, 0.005, 1.85, mask2) as lat:
        lat.set_cells(f0)
        lat.run(7)
        want, got = _both(lat)
    np.testing.assert_allclose(got, want, rtol=RTOL)
    rho = np.float64(f0.sum())
    assert abs(float(got[1]) - rho) <= 1e-5 * rho               # and it is the right number: mass is conserved


def test_noalloc_diagnostics_reject_more_rows_than_the_fixed_scratch_holds(lbm):
    ny = 16384 + 4                                              # LBM_NOALLOC_MAX_ROWS + 4 rows on one GPU
    with lbm.Lattice(4, ny, 2, 0.1, 0.005, 1.85, np.zeros((ny, 4), np.int32)) as lat:
        lat.run(2)
        assert np.isfinite(lat.total_density())                 # the allocating call still works
        with pytest.raises(lbm.LbmError, match="LBM_NOALLOC_MAX_ROWS"):
            lat.diagnostics_noalloc()


def test_noalloc_diagnostics_on_every_gpu_of_a_row_decomposition(lbm):
    """Each GPU holds its own copy of the array; its address is looked up once per device with that device current."""
    n = min(lbm.device_count(), 4)
    if n < 2:
        pytest.skip("needs 2 GPUs")
    p, mask, lines = load_grid("128x256")
    with lbm.Lattice(p["nx"], p["ny"], 120, p["density"], p["accel"], p["omega"], mask, obstacle_lines=lines, ngpus=n) as lat:
        lat.run(120)
        want, got = _both(lat)
        want_b, got_b = _both(lat)
    np.testing.assert_allclose(got, want, rtol=RTOL)
    np.testing.assert_allclose(got_b, want_b, rtol=RTOL)

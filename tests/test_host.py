"""CPU tests (-m "not gpu"): the C ABI library loads and exports every symbol include/lbm_b200.h
declares, the host-side planner is consistent, the host driver honours the reference's CLI / error
conventions, and the product refuses to run without a GPU (no CPU fallback)."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "d2q9-bgk")


@pytest.fixture(scope="module", autouse=True)
def built():
    if not (os.path.exists(EXE) and os.path.exists(os.path.join(ROOT, "lattice-boltzmann_b200", "liblbm_b200.so"))):
        subprocess.run(["make", "all"], cwd=ROOT, check=True, stdout=subprocess.DEVNULL)


def test_library_exports_every_declared_symbol(lbm):
    with open(os.path.join(ROOT, "include", "lbm_b200.h")) as fh:
        header = fh.read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(lbm_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 25
    lib = lbm.load_library()
    for name in declared:
        assert hasattr(lib, name), f"liblbm_b200.so does not export {name}"
    assert declared == set(lbm.SYMBOLS), declared ^ set(lbm.SYMBOLS)
    assert lib.lbm_abi_version() == 1


def test_no_cpu_fallback(lbm):
    if lbm.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(lbm.LbmError, match="no CPU fallback"):
        lbm.Lattice(128, 128, 10, 0.1, 0.005, 1.85, np.zeros((128, 128), np.int32))


def test_oracle_is_not_reachable_from_the_product():
    """Nothing shipped may import, link or call anything under oracle/."""
    for rel in ("lattice-boltzmann_b200/__init__.py", "lattice-boltzmann_b200/csrc/lbm_api.cu",
                "lattice-boltzmann_b200/csrc/lbm_kernels.cuh", "lattice-boltzmann_b200/csrc/lbm_device.cuh",
                "lattice-boltzmann_b200/csrc/lbm_plan.cpp", "src/d2q9-bgk.c"):
        with open(os.path.join(ROOT, rel)) as fh:
            text = fh.read()
        assert "oracle" not in text.lower(), f"{rel} mentions the oracle"
    out = subprocess.run(["ldd", EXE], capture_output=True, text=True).stdout
    assert "oracle" not in out and "liblbm_b200" in out


@pytest.mark.parametrize("ny,size", [(1024, 1), (1024, 3), (1024, 8), (128, 7), (5, 5), (256, 6)])
def test_allocate_rows_follows_reference_rule(lbm, ny, size):
    """d2q9-bgk.c:484-496: ny/size rows each, the first ny%size ranks one more, contiguous."""
    start = 0
    for r in range(size):
        s, n = lbm.allocate_rows(ny, size, r)
        assert s == start and n == ny // size + (1 if r < ny % size else 0)
        start += n
    assert start == ny


@pytest.mark.parametrize("nx,ny", [(128, 128), (128, 256), (1024, 1024), (16384, 1024), (132, 40), (4, 3), (2048, 7)])
@pytest.mark.parametrize("size", [1, 2, 3, 8])
def test_planner_selfcheck(lbm, nx, ny, size):
    if size > ny:
        pytest.skip("more slabs than rows")
    lib = lbm.load_library()
    for ctas in (1, 5, 148, 296):
        assert lib.lbm_plan_selftest(nx, ny, size, ctas) == 0, lib.lbm_last_error().decode()


@pytest.mark.parametrize("nx,ny,size", [(128, 128, 1), (132, 40, 3), (4, 3, 3), (1028, 9, 2), (2048, 1031, 8), (16384, 1024, 8)])
def test_host_setup_selfcheck(lbm, nx, ny, size):
    """The threaded obstacle count and bit-pack of lbm_create against plain loops (no device needed),
    with 1, 3 and the default number of host threads."""
    lib = lbm.load_library()
    for threads in ("1", "3", None):
        if threads is None:
            os.environ.pop("LBM_HOST_THREADS", None)
        else:
            os.environ["LBM_HOST_THREADS"] = threads
        try:
            assert lib.lbm_host_selftest(nx, ny, size, 7 + nx) == 0, lib.lbm_last_error().decode()
        finally:
            os.environ.pop("LBM_HOST_THREADS", None)


def test_planner_selfcheck_random_shapes(lbm):
    """Property test of the host planner over random lattices, slab counts and CTA budgets (hypothesis):
    every task scheduled once, interior tasks self-contained, dependencies complete, edge tasks first."""
    from hypothesis import given, settings, strategies as st
    lib = lbm.load_library()

    @settings(max_examples=1000, deadline=None, derandomize=True)
    @given(st.integers(1, 700), st.integers(3, 260), st.integers(1, 8), st.integers(1, 420), st.sampled_from([16, 24, 40, 64]))
    def check(nx4, ny, size, ctas, edge_weight):
        size = min(size, ny)
        os.environ["LBM_EDGE_WEIGHT_16"] = str(edge_weight)      # weight of a slab-edge task when the ranges are cut
        assert lib.lbm_plan_selftest(4 * nx4, ny, size, ctas) == 0, (4 * nx4, ny, size, ctas, lib.lbm_last_error().decode())

    try:
        check()
    finally:
        os.environ.pop("LBM_EDGE_WEIGHT_16", None)


def test_exact_text_formatter_matches_printf(tmp_path):
    """src/fmt_e12.h (the integer-arithmetic "%.12E" / "%d" used for final_state.dat and av_vels.dat) against
    snprintf byte for byte: special values, every binade's edges, values that round across a power of ten,
    4 M random bit patterns over all exponents and 4 M in the LBM range (tests/fmt_check.c)."""
    exe = tmp_path / "fmt_check"
    subprocess.run(["gcc", "-O2", "-std=gnu99", "-Wall", os.path.join(ROOT, "tests", "fmt_check.c"), "-o", str(exe)], check=True)
    r = subprocess.run([str(exe), "4"], capture_output=True, text=True)
    assert r.returncode == 0 and " 0 mismatches" in r.stdout, r.stdout[-2000:]


@pytest.mark.parametrize("nx,ny,threads", [(260, 37, "8"), (4, 3, "1"), (16384, 9, "5"), (132, 700, "3")])
def test_driver_writers_byte_identical_to_reference_formats(tmp_path, nx, ny, threads):
    """The host driver's own write_final_state() / write_av_vels() (compiled in from src/d2q9-bgk.c) on a
    synthetic lattice with zeros, negative zeros, tiny and ordinary values: files byte-identical to the
    reference's fprintf formats (d2q9-bgk.c:739, 747), whatever the thread count (tests/cli_writer_check.c)."""
    exe = tmp_path / "writer_check"
    pkg = os.path.join(ROOT, "lattice-boltzmann_b200")
    subprocess.run(["gcc", "-std=c99", "-Wall", "-O2", "-fopenmp", os.path.join(ROOT, "tests", "cli_writer_check.c"),
                    "-L" + pkg, "-llbm_b200", "-Wl,-rpath," + pkg, "-lm", "-o", str(exe)], check=True)
    r = subprocess.run([str(exe), str(nx), str(ny)], capture_output=True, text=True, cwd=tmp_path,
                       env=dict(os.environ, OMP_NUM_THREADS=threads))
    assert r.returncode == 0 and "writers identical" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def run_exe(*args, cwd=None):
    return subprocess.run([EXE, *args], capture_output=True, text=True, cwd=cwd)


def test_driver_usage_and_die_conventions(tmp_path):
    """Same messages / exit status as the reference (d2q9-bgk.c:166-168, 516-528, 621-633, 754-764)."""
    r = run_exe()
    assert r.returncode == 1 and r.stderr == f"Usage: {EXE} <paramfile> <obstaclefile>\n"
    r = run_exe("missing.params", "x.dat")
    assert r.returncode == 1 and "could not open input parameter file: missing.params" in r.stderr
    assert re.match(r"Error at line \d+ of file .*d2q9-bgk\.c:\n", r.stderr)
    good = os.path.join(ROOT, "inputs", "input_128x128.params")
    short = tmp_path / "short.params"
    short.write_text("128\n128\n")
    r = run_exe(str(short), "x.dat")
    assert r.returncode == 1 and "could not read param file: maxIters" in r.stderr
    r = run_exe(good, str(tmp_path / "nope.dat"))
    assert r.returncode == 1 and "could not open input obstacles file:" in r.stderr
    for text, msg in (("1 2\n", "expected 3 values per line in obstacle file"), ("128 0 1\n", "obstacle x-coord out of range"),
                      ("0 -1 1\n", "obstacle y-coord out of range"), ("3 3 2\n", "obstacle blocked value should be 1")):
        bad = tmp_path / "bad.dat"
        bad.write_text(text)
        r = run_exe(good, str(bad))
        assert r.returncode == 1 and msg in r.stderr, (text, r.stderr)


def test_checker_restatement_matches_reference_checker(tmp_path, golden):
    """tools/check.py prints and decides exactly like check/check.py (run where the reference exists)."""
    ref_checker = "/root/reference/check/check.py"
    if not os.path.exists(ref_checker):
        pytest.skip("reference tree not present on this box")
    import oracle_lib
    g = golden("ref_short_128x128_1000")
    ny, nx = g["pressure"].shape
    (tmp_path / "final_state.dat").write_text(oracle_lib.format_final_state(g["ux"], g["uy"], g["u"], g["pressure"], g["obstacles"]))
    (tmp_path / "av_vels.dat").write_text(oracle_lib.format_av_vels(g["av_vels"]))
    bad_p = g["pressure"].copy()
    bad_p[5, 5] *= 1.05
    (tmp_path / "final_state_bad.dat").write_text(oracle_lib.format_final_state(g["ux"], g["uy"], g["u"], bad_p, g["obstacles"]))
    for fs, expect in (("final_state.dat", 0), ("final_state_bad.dat", 1)):
        args = ["--ref-av-vels-file", str(tmp_path / "av_vels.dat"), "--ref-final-state-file", str(tmp_path / "final_state.dat"),
                "--av-vels-file", str(tmp_path / "av_vels.dat"), "--final-state-file", str(tmp_path / fs)]
        a = subprocess.run([sys.executable, ref_checker, *args], capture_output=True, text=True)
        b = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "check.py"), *args], capture_output=True, text=True)
        assert a.returncode == b.returncode == expect
        assert a.stdout == b.stdout


@pytest.mark.parametrize("lib", ["liblbm_b200.so", "liblbm_b200_checked.so"])
def test_tagged_record_accesses_are_never_narrowed_in_sass(lib):
    """A tagged record {v,v,v,tag} must be read and written by ONE 16-byte access, or a reader can pair a new tag with
    an old payload.  ptxas does not treat ld.relaxed.sys.b128 as indivisible: where only some words are used it
    narrowed the load, and in the self-checking build it split it into two 32-bit loads (payload first, tag second),
    which broke engine 3 there.  The accessors are .volatile now (lbm_device.cuh); this test reads the SASS of every
    timestep kernel of every build: all system-scope strong accesses - which only the records use - are 128 bits wide."""
    import shutil
    path = os.path.join(ROOT, "lattice-boltzmann_b200", lib)
    if not os.path.exists(path) or not shutil.which("cuobjdump"):
        pytest.skip("needs the built library and cuobjdump")
    sass = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True, check=True).stdout
    fn, per_fn = None, {}
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            fn = m.group(1)
            continue
        m = re.search(r"\b(LDG|STG)\.E((?:\.[A-Z0-9]+)*)\.STRONG\.SYS\b", line)
        if m and fn and re.search(r"lbm_(persistent|dataflow|row|step)_kernel", fn):
            per_fn.setdefault(fn, []).append(m.group(1) + m.group(2))
    assert per_fn, "no timestep kernels found in the SASS"
    total = 0
    for fn, ops in per_fn.items():
        narrow = [o for o in ops if not o.endswith(".128")]
        assert not narrow, f"{fn}: narrowed record accesses {sorted(set(narrow))}"
        total += len(ops)
    assert total > 100

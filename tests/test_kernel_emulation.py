"""The row-stream kernel skeleton (or-tools_b200/csrc/pdlp_skeletons.cuh) run on CPU.

tests/kernel_emu/ compiles the kernel source itself as host C++ against a shim (a CTA = 256 host
threads, __syncthreads / shuffles = barriers, cp.async landing either at issue or at the wait) and
runs it on the real work-block partition (pdlp_b200.work_blocks).  What this pins down without a
GPU: the index arithmetic of the staging and reduction phases (bank-padded positions, threads per
row, prefetch distances, the long-row path, ragged and empty rows, the per-thread tail positions),
the placement of waits and barriers, and that the PDLP_ROWSTREAM_LEAN variant computes bit-identical
results to the default path.  It says nothing about speed or about the hardware memory model; the
GPU parity tests (tests/test_gpu_solve.py) remain the check of the shipped binary."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

import pdlp_b200

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EMU = os.path.join(ROOT, "tests", "kernel_emu")
BUILD = os.path.join(ROOT, "tests", "_build")
CSRC = os.path.join(ROOT, "or-tools_b200", "csrc")


def _build(lean):
    os.makedirs(BUILD, exist_ok=True)
    out = os.path.join(BUILD, f"rowstream_emu_lean{lean}.so")
    sources = [os.path.join(EMU, "rowstream_emu.cc"), os.path.join(EMU, "cuda_emu.h"), os.path.join(CSRC, "pdlp_skeletons.cuh")]
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(s) for s in sources):
        subprocess.run(["g++", "-std=c++20", "-O1", "-pthread", "-shared", "-fPIC", "-Wno-unknown-pragmas", "-Wno-attributes",
                        f"-DPDLP_ROWSTREAM_LEAN={lean}", "-I" + EMU, "-I" + CSRC, sources[0], "-o", out], check=True)
    lib = ctypes.CDLL(out)
    assert lib.emu_is_lean() == lean
    return lib


@pytest.fixture(scope="module")
def libs():
    return {0: _build(0), 1: _build(1)}


def make_matrix(lengths, cols, rng, integer=True):
    lengths = np.asarray(lengths, dtype=np.int64)
    row_ptr = np.concatenate([[0], np.cumsum(lengths)]).astype(np.int32)
    nnz = int(row_ptr[-1])
    idx = rng.integers(0, cols, max(nnz, 1)).astype(np.int32)
    if integer:  # every partial sum is exact, whatever the order
        val = rng.integers(-3, 4, max(nnz, 1)).astype(np.float64)
        x = rng.integers(-4, 5, cols).astype(np.float64)
        bias = rng.integers(-9, 10, len(lengths)).astype(np.float64)
    else:
        val = rng.standard_normal(max(nnz, 1))
        x = rng.standard_normal(cols)
        bias = rng.standard_normal(len(lengths))
    return row_ptr, idx, val, x, bias


def run(lib, op, mat, grid, defer=0, ulps=0):
    row_ptr, idx, val, x, bias = mat
    rows = len(row_ptr) - 1
    starts, _, _ = pdlp_b200.work_blocks(row_ptr)
    starts = np.ascontiguousarray(starts, dtype=np.int32)
    nb = len(starts) - 1
    out = np.full(rows, np.nan)
    totals = np.zeros(2)
    p = lambda a: ctypes.c_void_p(a.ctypes.data)
    rc = lib.emu_rowstream(op, p(row_ptr), p(idx), p(val), p(x), p(bias), rows, p(starts), nb, min(grid, nb), defer, ulps, p(out), p(totals))
    assert rc == 0
    return out, totals


def reference(op, mat):
    row_ptr, idx, val, x, bias = mat
    rows = len(row_ptr) - 1
    prod = val * x[idx]
    out = np.zeros(rows)
    for i in range(rows):
        seg = prod[row_ptr[i]:row_ptr[i + 1]]
        out[i] = seg.sum() if op == 0 else (np.abs(seg).max() if seg.size else 0.0)
    if op == 0:
        return out + bias, np.array([out.sum(), np.abs(out).max() if rows else -np.inf])
    return out, np.array([out.max() if rows else -np.inf, 0.0])


SHAPES = {
    "uniform16": lambda rng: np.full(700, 16),                # full blocks, 2 threads per row
    "uniform3": lambda rng: np.full(2100, 3),                 # odd length: no bank pad
    "uniform64": lambda rng: np.full(100, 64),                # 8 threads per row
    "poisson10": lambda rng: rng.poisson(10, 900),            # ragged rows, blocks a few short of full
    "empty_rows": lambda rng: rng.integers(0, 2, 5000),       # 1024-row blocks, one thread per row, tail loop
    "wide_rows": lambda rng: rng.integers(300, 2048, 12),     # 1-6 rows per block: 32 threads per row
    "long_rows": lambda rng: np.array([5, 0, 5000, 7, 2048, 2049, 1, 9000, 3, 3]),  # CTA-wide path between staged blocks
    "short_tail": lambda rng: np.concatenate([np.full(256, 8), [5]]),               # last block of 5 nonzeros
    "one_nonzero": lambda rng: np.array([1]),
    "ends_at_1700": lambda rng: np.full(170, 10),             # last block just below the all-but-last-position-full rule
    "ends_at_1800": lambda rng: np.full(180, 10),             # ... and just above it: only the tail position is predicated
}


@pytest.mark.parametrize("lean", [0, 1])
@pytest.mark.parametrize("shape", sorted(SHAPES))
def test_rowstream_matches_a_serial_product_exactly(libs, shape, lean):
    rng = np.random.default_rng(17)
    mat = make_matrix(SHAPES[shape](rng), 333, rng)
    for op in (0, 1):
        want_out, want_tot = reference(op, mat)
        for grid, defer in ((1, 0), (3, 1)):
            out, tot = run(libs[lean], op, mat, grid, defer)
            np.testing.assert_array_equal(out, want_out, err_msg=f"op {op} grid {grid} defer {defer}")
            np.testing.assert_array_equal(tot[:2 if op == 0 else 1], want_tot[:2 if op == 0 else 1])


@pytest.mark.parametrize("shape", ["uniform16", "poisson10", "wide_rows", "long_rows", "empty_rows"])
def test_lean_staging_is_bit_identical_to_the_default_path(libs, shape):
    rng = np.random.default_rng(23)
    mat = make_matrix(SHAPES[shape](rng), 1000, rng, integer=False)
    base_out, base_tot = run(libs[0], 0, mat, 2, 0)
    for defer in (0, 1):
        out, tot = run(libs[1], 0, mat, 2, defer)
        np.testing.assert_array_equal(out, base_out)
        np.testing.assert_array_equal(tot, base_tot)


@pytest.mark.parametrize("ulps", [-2, 2])
def test_lean_bank_pad_rule_tolerates_the_fast_division_error(libs, ulps):
    # __fdividef is accurate to 2 ulp; the pad rule must pick the same layout (results are layout
    # independent, so this also checks that a different pad would still be correct).
    rng = np.random.default_rng(29)
    for shape in ("uniform16", "uniform3", "poisson10"):
        mat = make_matrix(SHAPES[shape](rng), 500, rng)
        want_out, _ = reference(0, mat)
        out, _ = run(libs[1], 0, mat, 2, 1, ulps)
        np.testing.assert_array_equal(out, want_out)


@pytest.mark.parametrize("count, grid", [(0, 1), (1, 1), (255, 1), (256, 2), (1000, 3), (5000, 7)])
def test_dense_skeleton_and_its_two_stage_reduction(libs, count, grid):
    rng = np.random.default_rng(count + grid)
    a = rng.integers(-5, 6, max(count, 1)).astype(np.float64)
    b0 = rng.integers(-5, 6, max(count, 1)).astype(np.float64)
    p = lambda v: ctypes.c_void_p(v.ctypes.data)
    lib = libs[0]
    lib.emu_map_reduce.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_double, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    b, totals = b0.copy(), np.zeros(3)
    assert lib.emu_map_reduce(p(a), p(b), count, 2.0, 1, grid, p(totals)) == 0
    want = (b0 + 2.0 * a)[:count]
    np.testing.assert_array_equal(b[:count], want)
    assert totals[0] == want.sum() and totals[2] == want.sum()
    assert totals[1] == (np.abs(want).max() if count else -np.inf)
    # Begin() == false: the device-side early-out leaves the data alone and reduces to the identities
    b, totals = b0.copy(), np.zeros(3)
    assert lib.emu_map_reduce(p(a), p(b), count, 2.0, 0, grid, p(totals)) == 0
    np.testing.assert_array_equal(b, b0)
    assert totals[0] == 0.0 and totals[1] == -np.inf and totals[2] == 0.0

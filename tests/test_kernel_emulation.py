"""The row-stream kernel skeleton (or-tools_b200/csrc/pdlp_skeletons.cuh) run on CPU.

tests/kernel_emu/ compiles the kernel source itself as host C++ against a shim (a CTA = 256 host
threads, __syncthreads / shuffles = barriers, cp.async landing either at issue or at the wait) and
runs it on the real work-block partition (pdlp_b200.work_blocks).  What this pins down without a
GPU: the index arithmetic of the staging and reduction phases (bank-padded positions, threads per
row, prefetch distances, the long-row path, ragged and empty rows, the per-thread tail positions),
and the placement of waits and barriers.  It says nothing about speed or about the hardware memory model; the
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


def _build():
    os.makedirs(BUILD, exist_ok=True)
    out = os.path.join(BUILD, "rowstream_emu.so")
    sources = [os.path.join(EMU, "rowstream_emu.cc"), os.path.join(EMU, "cuda_emu.h"), os.path.join(CSRC, "pdlp_skeletons.cuh")]
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(s) for s in sources):
        subprocess.run(["g++", "-std=c++20", "-O1", "-pthread", "-shared", "-fPIC", "-DPDLP_EMU_FIBERS", "-Wno-unknown-pragmas", "-Wno-attributes",
                        "-I" + EMU, "-I" + CSRC, sources[0], "-o", out], check=True)
    return ctypes.CDLL(out)


@pytest.fixture(scope="module")
def libs():
    return {0: _build()}


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


@pytest.mark.parametrize("shape", sorted(SHAPES))
def test_rowstream_matches_a_serial_product_exactly(libs, shape, lean=0):
    rng = np.random.default_rng(17)
    mat = make_matrix(SHAPES[shape](rng), 333, rng)
    for op in (0, 1):
        want_out, want_tot = reference(op, mat)
        for grid, defer in ((1, 0), (3, 1)):
            out, tot = run(libs[lean], op, mat, grid, defer)
            np.testing.assert_array_equal(out, want_out, err_msg=f"op {op} grid {grid} defer {defer}")
            np.testing.assert_array_equal(tot[:2 if op == 0 else 1], want_tot[:2 if op == 0 else 1])


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


# ---- the adaptive trial step (K1 - K4 + averaging) against a restatement of the reference ----

def _build_step():
    out = os.path.join(BUILD, "step_emu.so")
    sources = [os.path.join(EMU, "step_emu.cc"), os.path.join(EMU, "cuda_emu.h"), os.path.join(CSRC, "pdlp_skeletons.cuh"),
               os.path.join(CSRC, "pdlp_step_kernels.cuh"), os.path.join(CSRC, "pdlp_step_state.h"), os.path.join(CSRC, "pdlp_peer.cuh"), os.path.join(CSRC, "pdlp_peer_view.h")]
    os.makedirs(BUILD, exist_ok=True)
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(s) for s in sources):
        subprocess.run(["g++", "-std=c++20", "-O1", "-pthread", "-shared", "-fPIC", "-DPDLP_EMU_FIBERS", "-Wno-unknown-pragmas", "-Wno-attributes",
                        "-I" + EMU, "-I" + CSRC, sources[0], "-o", out], check=True)
    return ctypes.CDLL(out)


@pytest.fixture(scope="module")
def step_libs():
    return {0: _build_step()}


def reference_adaptive_trials(K, c, q, vlb, vub, clb, cub, x, y, step, w, trials, constant=False):
    """Solver::TakeAdaptiveStep (primal_dual_hybrid_gradient.cc:2295-2376) one trial at a time, with
    ComputeNextPrimalSolution (:1659-1705), ComputeNextDualSolution (:1707-1748), ComputeMovement
    (:1750-1759), ComputeNonlinearity (:1761-1771) and ShardedWeightedAverage::Add
    (sharded_optimization_utils.cc:54-66).  constant=True: TakeConstantSizeStep (:2378-2406), every
    step accepted at the same size.  Dense numpy; K is a scipy matrix."""
    x, y = x.copy(), y.copy()
    kty = K.T @ y
    avg_x, avg_y, weight_sum = np.zeros_like(x), np.zeros_like(y), 0.0
    completed = rejected = inner = 0
    movement = nonlinearity = 0.0
    for _ in range(trials):
        tau, sigma = step / w, step * w
        t = x - tau * (c - kty)
        if q is not None:
            t = t / (1.0 + tau * q)
        xn = np.minimum(np.maximum(t, vlb), vub)
        dx = xn - x
        temp = y - sigma * (K @ (xn + dx))
        yn = np.maximum(np.minimum(0.0, temp + sigma * cub), temp + sigma * clb)
        dy = yn - y
        movement = 0.5 * w * (dx @ dx) + 0.5 / w * (dy @ dy)
        assert 0.0 < movement < 1e100
        ktyn = K.T @ yn
        nonlinearity = -(dx @ (ktyn - kty))
        limit = movement / nonlinearity if nonlinearity > 0 else np.inf
        accepted = constant or step <= limit
        total = rejected + inner + completed + 1
        if accepted:
            x, y, kty = xn, yn, ktyn
            ratio = step / (weight_sum + step)
            avg_x += ratio * (x - avg_x)
            avg_y += ratio * (y - avg_y)
            weight_sum += step
        first = limit if np.isinf(limit) else (1 - (total + 1.0) ** -0.3) * limit
        second = (1 + (total + 1.0) ** -0.6) * step
        if not constant:
            step = min(first, second)
        if accepted:
            rejected += inner  # num_rejected_steps_ += inner_iterations - 1 after the loop
            inner = 0
            completed += 1
        else:
            inner += 1
    return dict(x=x, y=y, avg_x=avg_x, avg_y=avg_y, step=step, completed=completed, rejected=rejected + inner, movement=movement,
                nonlinearity=nonlinearity, weight_sum=weight_sum)


def emulated_adaptive_trials(lib, K, c, q, vlb, vub, clb, cub, x, y, step, w, trials, max_grid=3, defer=0, constant=False, interrupt_at=-1):
    import scipy.sparse as sp
    csr, csc = sp.csr_matrix(K), sp.csc_matrix(K)
    csr.sort_indices(); csc.sort_indices()
    m, n = K.shape
    i32 = lambda a: np.ascontiguousarray(a, dtype=np.int32)
    f64 = lambda a: np.ascontiguousarray(a, dtype=np.float64)
    arrays = dict(rp=i32(csr.indptr), ri=i32(csr.indices), rv=f64(csr.data), cp=i32(csc.indptr), ci=i32(csc.indices), cv=f64(csc.data))
    rb, _, _ = pdlp_b200.work_blocks(arrays["rp"])
    cb, _, _ = pdlp_b200.work_blocks(arrays["cp"])
    rb, cb = i32(rb), i32(cb)
    xo, yo, ax, ay, st = f64(x).copy(), f64(y).copy(), np.zeros(n), np.zeros(m), np.zeros(9)
    p = lambda a: ctypes.c_void_p(a.ctypes.data) if a is not None else ctypes.c_void_p(None)
    vecs = [f64(c), None if q is None else f64(q), f64(vlb), f64(vub), f64(clb), f64(cub)]
    lib.emu_set_interrupt_at(int(interrupt_at))  # the flag is raised before this trial of the batch (-1: no flag)
    rc = lib.emu_adaptive_trials(n, m, p(arrays["rp"]), p(arrays["ri"]), p(arrays["rv"]), p(rb), len(rb) - 1, p(arrays["cp"]), p(arrays["ci"]), p(arrays["cv"]),
                                 p(cb), len(cb) - 1, *[p(v) for v in vecs], p(xo), p(yo), p(ax), p(ay), ctypes.c_double(step), ctypes.c_double(w), trials, int(constant),
                                 max_grid, defer, p(st))
    lib.emu_set_interrupt_at(-1)
    assert rc == 0
    return dict(x=xo, y=yo, avg_x=ax, avg_y=ay, step=st[0], completed=int(st[1]), rejected=int(st[2]), movement=st[3], nonlinearity=st[4], weight_sum=st[5],
                active=int(st[6]), total_trials=int(st[7]), stopped_by_interrupt=int(st[8]))


def step_problem(seed, qp):
    import scipy.sparse as sp
    rng = np.random.default_rng(seed)
    n, m = 700, 420
    K = sp.random(m, n, density=0.02, random_state=seed, data_rvs=lambda k: rng.standard_normal(k)).tocsr()
    c = rng.standard_normal(n)
    q = rng.uniform(0.0, 2.0, n) if qp else None
    vlb = np.where(rng.random(n) < 0.3, -np.inf, rng.uniform(-2, 0, n))
    vub = np.where(rng.random(n) < 0.3, np.inf, rng.uniform(0, 2, n))
    mid = rng.standard_normal(m)
    clb = np.where(rng.random(m) < 0.3, -np.inf, mid - rng.uniform(0, 1, m))
    cub = np.where(rng.random(m) < 0.3, np.inf, mid + rng.uniform(0, 1, m))
    eq = rng.random(m) < 0.2
    clb[eq] = cub[eq] = mid[eq]
    x0 = np.minimum(np.maximum(rng.standard_normal(n), vlb), vub)
    y0 = rng.standard_normal(m) * 0.1
    y0 = np.where(np.isinf(clb), np.minimum(y0, 0.0), y0)
    y0 = np.where(np.isinf(cub), np.maximum(y0, 0.0), y0)
    return K, c, q, vlb, vub, clb, cub, x0, y0


@pytest.mark.parametrize("defer", [0, 1])
@pytest.mark.parametrize("qp", [False, True])
def test_adaptive_trial_steps_follow_the_reference_step_rule(step_libs, qp, defer):
    prob = step_problem(2 if not qp else 6, qp)
    # a first step far too long: the rule must reject, shrink, then accept and grow again
    step, w, trials = 50.0, 0.2, 14
    want = reference_adaptive_trials(*prob, step, w, trials)
    assert want["rejected"] >= 2 and want["completed"] >= 8
    got = emulated_adaptive_trials(step_libs[0], *prob, step, w, trials, defer=defer)
    assert (got["completed"], got["rejected"]) == (want["completed"], want["rejected"])
    assert got["active"] == 1 and got["total_trials"] == trials
    for key in ("x", "y", "avg_x", "avg_y"):
        np.testing.assert_allclose(got[key], want[key], rtol=1e-11, atol=1e-12, err_msg=key)
    for key in ("step", "movement", "nonlinearity", "weight_sum"):
        np.testing.assert_allclose(got[key], want[key], rtol=1e-10, err_msg=key)


def test_decision_kernel_stops_at_the_first_accepted_step_after_the_interrupt(step_libs):
    """termination.cc:178 tests interrupt_solve between iterations: a flag raised during an iteration's rejected trials
    stops the loop when that iteration completes, not earlier and not a whole polling interval later."""
    prob = step_problem(2, False)
    step, w, trials = 50.0, 0.2, 14
    free = [reference_adaptive_trials(*prob, step, w, k) for k in range(trials + 1)]
    accepted_at = [k for k in range(trials) if free[k + 1]["completed"] > free[k]["completed"]]  # trial k was accepted
    rejected_at = [k for k in range(trials) if k not in accepted_at]
    assert rejected_at and accepted_at
    for raise_before in (rejected_at[0], accepted_at[2]):
        got = emulated_adaptive_trials(step_libs[0], *prob, step, w, trials, interrupt_at=raise_before)
        stop = min(k for k in accepted_at if k >= raise_before)  # the trial that completes the running iteration
        assert got["stopped_by_interrupt"] == 1 and got["active"] == 0
        assert got["total_trials"] == stop + 1, (got["total_trials"], stop)
        want = free[stop + 1]
        assert got["completed"] == want["completed"]
        for key in ("x", "y", "avg_x", "avg_y"):
            np.testing.assert_allclose(got[key], want[key], rtol=1e-11, atol=1e-12, err_msg=key)
    # No flag: nothing stops.
    got = emulated_adaptive_trials(step_libs[0], *prob, step, w, trials)
    assert got["stopped_by_interrupt"] == 0 and got["active"] == 1 and got["total_trials"] == trials


def test_constant_size_steps_follow_the_reference(step_libs):
    prob = step_problem(3, False)
    step, w, trials = 0.05, 0.7, 6
    want = reference_adaptive_trials(*prob, step, w, trials, constant=True)
    got = emulated_adaptive_trials(step_libs[0], *prob, step, w, trials, constant=True)
    assert (got["completed"], got["rejected"], got["step"]) == (trials, 0, step)
    for key in ("x", "y", "avg_x", "avg_y"):
        np.testing.assert_allclose(got[key], want[key], rtol=1e-11, atol=1e-12, err_msg=key)
    np.testing.assert_allclose(got["weight_sum"], want["weight_sum"], rtol=1e-12)


def reference_malitsky_pock_trials(K, c, q, vlb, vub, clb, cub, x, y, step, w, trials, downscaling=0.7, contraction=0.99, interpolation=1.0):
    """Solver::TakeMalitskyPockStep (primal_dual_hybrid_gradient.cc:2211-2293), one dual trial at a time."""
    x, y = x.copy(), y.copy()
    kty = K.T @ y
    avg_x, avg_y, wx, wy = np.zeros_like(x), np.zeros_like(y), 0.0, 0.0
    completed = rejected = 0
    ratio = 1.0
    new_tau = None
    for _ in range(trials):
        tau = step / w
        if new_tau is None:  # a new outer iteration
            t = x - tau * (c - kty)
            if q is not None:
                t = t / (1.0 + tau * q)
            xn = np.minimum(np.maximum(t, vlb), vub)
            dx = xn - x
            new_tau = tau * (1 + interpolation * (np.sqrt(1 + ratio) - 1))
        theta = new_tau / tau
        sigma = w * w * new_tau
        temp = y - sigma * (K @ (xn + theta * dx))
        yn = np.maximum(np.minimum(0.0, temp + sigma * cub), temp + sigma * clb)
        dy = yn - y
        ktyn = K.T @ yn
        if w * new_tau * np.linalg.norm(ktyn - kty) <= contraction * np.linalg.norm(dy):
            step, ratio = new_tau * w, theta
            if wx == 0.0:
                weight = new_tau * theta
                avg_x += weight / (wx + weight) * (x - avg_x)
                wx += weight
            x, y, kty = xn, yn, ktyn
            avg_x += new_tau / (wx + new_tau) * (x - avg_x)
            wx += new_tau
            avg_y += new_tau / (wy + new_tau) * (y - avg_y)
            wy += new_tau
            completed += 1
            new_tau = None
        else:
            new_tau = downscaling * new_tau
            rejected += 1
    return dict(x=x, y=y, avg_x=avg_x, avg_y=avg_y, step=step, completed=completed, rejected=rejected, wx=wx, wy=wy, ratio=ratio)


def test_malitsky_pock_steps_follow_the_reference(step_libs):
    import scipy.sparse as sp
    K, c, q, vlb, vub, clb, cub, x0, y0 = prob = step_problem(4, False)
    step, w, trials = 0.08, 0.8, 16
    want = reference_malitsky_pock_trials(*prob, step, w, trials)
    assert want["rejected"] >= 2 and want["completed"] >= 6
    lib = step_libs[0]
    csr, csc = sp.csr_matrix(K), sp.csc_matrix(K)
    csr.sort_indices(); csc.sort_indices()
    m, n = K.shape
    i32 = lambda a: np.ascontiguousarray(a, dtype=np.int32)
    f64 = lambda a: np.ascontiguousarray(a, dtype=np.float64)
    rp, ri, rv, cp, ci, cv = i32(csr.indptr), i32(csr.indices), f64(csr.data), i32(csc.indptr), i32(csc.indices), f64(csc.data)
    rb, cb = i32(pdlp_b200.work_blocks(rp)[0]), i32(pdlp_b200.work_blocks(cp)[0])
    xo, yo, ax, ay, st = f64(x0).copy(), f64(y0).copy(), np.zeros(n), np.zeros(m), np.zeros(10)
    params = np.array([0.7, 0.99, 1.0])
    p = lambda a: ctypes.c_void_p(a.ctypes.data) if a is not None else ctypes.c_void_p(None)
    rc = lib.emu_malitsky_pock_trials(n, m, p(rp), p(ri), p(rv), p(rb), len(rb) - 1, p(cp), p(ci), p(cv), p(cb), len(cb) - 1, p(f64(c)), p(None), p(f64(vlb)),
                                      p(f64(vub)), p(f64(clb)), p(f64(cub)), p(xo), p(yo), p(ax), p(ay), ctypes.c_double(step), ctypes.c_double(w), p(params),
                                      trials, 3, p(st))
    assert rc == 0
    assert (int(st[1]), int(st[2])) == (want["completed"], want["rejected"])
    for got, key in ((xo, "x"), (yo, "y"), (ax, "avg_x"), (ay, "avg_y")):
        np.testing.assert_allclose(got, want[key], rtol=1e-11, atol=1e-12, err_msg=key)
    np.testing.assert_allclose([st[0], st[5], st[8], st[9]], [want["step"], want["wx"], want["wy"], want["ratio"]], rtol=1e-12)


# ---- the column-sharded dual step on emulated ranks ----

@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_dual_step_over_emulated_ranks_matches_one_rank(step_libs, world):
    """K2 of a column-sharded solve with the real host-side sharding: every rank's block
    holds the shared constraints first and its private ones after; phase 1 stores the shared rows'
    partial sums, the test adds them in rank order (the all-reduce), phase 2 finishes private rows
    inside the row-stream kernel and shared rows through FinishRowOp.  Gathered y' and the summed
    |dy|^2 must equal the unsharded step."""
    import scipy.sparse as sp
    rng = np.random.default_rng(40 + world)
    m, n = 900, 1600
    # banded columns (most constraints private to one rank) plus a few dense rows that every rank touches
    rows = np.concatenate([(np.arange(n * 4) // 4 * m // n + rng.integers(0, 30, n * 4)) % m, rng.integers(0, 12, 800)])
    cols = np.concatenate([np.arange(n * 4) // 4, rng.integers(0, n, 800)])
    K = sp.csc_matrix(sp.coo_matrix((rng.standard_normal(rows.size), (rows, cols)), shape=(m, n)))
    K.sum_duplicates()
    xbar, y = rng.standard_normal(n), rng.standard_normal(m) * 0.3
    mid = rng.standard_normal(m)
    clb = np.where(rng.random(m) < 0.3, -np.inf, mid - 0.5)
    cub = np.where(rng.random(m) < 0.3, np.inf, mid + 0.5)
    step, w = 0.4, 1.7
    sigma = step * w
    temp = y - sigma * (K @ xbar)
    want = np.maximum(np.minimum(0.0, temp + sigma * cub), temp + sigma * clb)

    starts = pdlp_b200.column_partition(K.indptr, world)
    sharing = pdlp_b200.sharding.RowSharing.from_matrix(K, starts)
    S = int(sharing.shared.size)
    assert 0 < S < m and all(p.size > 0 for p in sharing.private)
    lib = step_libs[0]
    p = lambda a: ctypes.c_void_p(a.ctypes.data)
    i32 = lambda a: np.ascontiguousarray(a, dtype=np.int32)
    ranks = []
    for r in range(world):
        local = sharing.local_rows(r)
        block = sp.csr_matrix(K[:, starts[r]:starts[r + 1]].tocsr()[local, :])
        block.sort_indices()
        ptr = i32(block.indptr)
        blocks, cut_block, _ = pdlp_b200.work_blocks(ptr, S if 0 < S < local.size else -1)
        if S >= local.size:
            cut_block = len(blocks) - 1
        ranks.append(dict(m=local.size, ptr=ptr, idx=i32(block.indices), val=np.ascontiguousarray(block.data), blocks=i32(blocks), cut=int(cut_block),
                          xbar=np.ascontiguousarray(xbar[starts[r]:starts[r + 1]]), y0=sharing.scatter(r, y), y1=np.zeros(local.size),
                          clb=sharing.scatter(r, clb), cub=sharing.scatter(r, cub)))
    sums = np.zeros(S)
    for k in ranks:  # phase 1 + all-reduce in rank order
        part = np.zeros(S)
        assert lib.emu_sharded_dual_partial(k["m"], p(k["ptr"]), p(k["idx"]), p(k["val"]), p(k["blocks"]), len(k["blocks"]) - 1, k["cut"], p(k["xbar"]), p(k["y0"]),
                                            p(k["y1"]), p(k["clb"]), p(k["cub"]), ctypes.c_double(step), ctypes.c_double(w), 3, p(part)) == 0
        sums += part
    np.testing.assert_allclose(sums, (K @ xbar)[sharing.shared], rtol=1e-11, atol=1e-12)
    total = 0.0
    for r, k in enumerate(ranks):
        dy2 = ctypes.c_double(0.0)
        assert lib.emu_sharded_dual_finish(k["m"], p(k["ptr"]), p(k["idx"]), p(k["val"]), p(k["blocks"]), len(k["blocks"]) - 1, k["cut"], S, p(k["xbar"]), p(k["y0"]),
                                           p(k["y1"]), p(k["clb"]), p(k["cub"]), ctypes.c_double(step), ctypes.c_double(w), 0 if r == 0 else S, 3, p(sums),
                                           ctypes.byref(dy2)) == 0
        total += dy2.value
        if r > 0:  # the shared part is replicated: bit-identical on every rank
            np.testing.assert_array_equal(k["y1"][:S], ranks[0]["y1"][:S])
    got = sharing.gather([k["y1"] for k in ranks])
    np.testing.assert_allclose(got, want, rtol=1e-11, atol=1e-12)
    np.testing.assert_allclose(total, np.sum((want - y) ** 2), rtol=1e-11)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_peer_memory_exchange_of_the_sharded_dual_step(step_libs, world):
    """The column-sharded dual step with the shared rows' sums exchanged through peer memory
    (pdlp_peer.cuh: PeerPartialRowOp pushes into every rank's region, peer_signal_rows_kernel publishes the
    epoch, PeerFinishRowOp waits and adds the N slots in rank order), all ranks emulated in one process.
    Three consecutive epochs with different operands: the parity double buffer must never serve an older
    epoch's sums, the replicated shared part must be bit-identical on every rank, and the result must equal
    the unsharded step."""
    import scipy.sparse as sp
    rng = np.random.default_rng(40 + world)
    m, n = 700, 1500
    K = sp.random(m, n, density=0.01, random_state=7, data_rvs=lambda k: rng.standard_normal(k)).tocsc()
    K = sp.csc_matrix(K)
    # a band of rows touched by one column block only (private) plus rows touched by all (shared)
    starts = pdlp_b200.column_partition(K.indptr, world)
    sharing = pdlp_b200.sharding.RowSharing.from_matrix(K, starts)
    S = int(sharing.shared.size)
    assert S > 0
    y = rng.standard_normal(m)
    mid = rng.standard_normal(m)
    clb = np.where(rng.random(m) < 0.3, -np.inf, mid - 0.5)
    cub = np.where(rng.random(m) < 0.3, np.inf, mid + 0.5)
    y = np.where(np.isinf(clb), np.minimum(y, 0.0), y)
    y = np.where(np.isinf(cub), np.maximum(y, 0.0), y)
    step, w = 0.4, 1.7
    sigma = step * w
    lib = step_libs[0]
    lib.emu_peer_region_bytes.restype = ctypes.c_longlong
    p = lambda a: ctypes.c_void_p(a.ctypes.data)
    i32 = lambda a: np.ascontiguousarray(a, dtype=np.int32)
    nbytes = lib.emu_peer_region_bytes(world, ctypes.c_longlong(S))
    regions = [np.zeros(nbytes // 8 + 1, dtype=np.float64) for _ in range(world)]
    table = (ctypes.c_void_p * world)(*[r.ctypes.data for r in regions])
    blocks_of = []
    for r in range(world):
        local = sharing.local_rows(r)
        block = sp.csr_matrix(K[:, starts[r]:starts[r + 1]].tocsr()[local, :])
        block.sort_indices()
        ptr = i32(block.indptr)
        blocks, cut_block, _ = pdlp_b200.work_blocks(ptr, S if 0 < S < local.size else -1)
        if S >= local.size:
            cut_block = len(blocks) - 1
        blocks_of.append(dict(m=local.size, ptr=ptr, idx=i32(block.indices), val=np.ascontiguousarray(block.data), blocks=i32(blocks), cut=int(cut_block),
                              y0=sharing.scatter(r, y), clb=sharing.scatter(r, clb), cub=sharing.scatter(r, cub)))
    for epoch in (1, 2, 3):
        xbar = rng.standard_normal(n) * epoch
        temp = y - sigma * (K @ xbar)
        want = np.maximum(np.minimum(0.0, temp + sigma * cub), temp + sigma * clb)
        ranks = [dict(k, xbar=np.ascontiguousarray(xbar[starts[r]:starts[r + 1]]), y1=np.zeros(k["m"])) for r, k in enumerate(blocks_of)]
        fused = epoch == 2  # the middle epoch through the one-launch form (PeerCutRowOp), the others as two launches
        total = 0.0
        for r, k in enumerate(ranks):
            if fused:
                part = ctypes.c_double(0.0)
                assert lib.emu_peer_dual_fused(k["m"], p(k["ptr"]), p(k["idx"]), p(k["val"]), p(k["blocks"]), len(k["blocks"]) - 1, S, p(k["xbar"]), p(k["y0"]),
                                               p(k["y1"]), p(k["clb"]), p(k["cub"]), ctypes.c_double(step), ctypes.c_double(w), 0 if r == 0 else S, 3, r, world,
                                               ctypes.c_longlong(S), table, ctypes.c_ulonglong(epoch), ctypes.byref(part)) == 0
                total += part.value
            else:
                assert lib.emu_peer_dual_partial(k["m"], p(k["ptr"]), p(k["idx"]), p(k["val"]), p(k["blocks"]), len(k["blocks"]) - 1, k["cut"], p(k["xbar"]), p(k["y0"]),
                                                 p(k["y1"]), p(k["clb"]), p(k["cub"]), ctypes.c_double(step), ctypes.c_double(w), 3, r, world, ctypes.c_longlong(S), table,
                                                 ctypes.c_ulonglong(epoch)) == 0
        for r, k in enumerate(ranks):
            dy2 = ctypes.c_double(0.0)
            nb = len(k["blocks"]) - 1
            assert lib.emu_peer_dual_finish(k["m"], p(k["ptr"]), p(k["idx"]), p(k["val"]), p(k["blocks"]), nb, nb if fused else k["cut"], S, p(k["xbar"]), p(k["y0"]),
                                            p(k["y1"]), p(k["clb"]), p(k["cub"]), ctypes.c_double(step), ctypes.c_double(w), 0 if r == 0 else S, 3, r, world,
                                            ctypes.c_longlong(S), table, ctypes.c_ulonglong(epoch), ctypes.byref(dy2)) == 0
            total += dy2.value
            if r > 0:
                np.testing.assert_array_equal(k["y1"][:S], ranks[0]["y1"][:S])
        got = sharing.gather([k["y1"] for k in ranks])
        np.testing.assert_allclose(got, want, rtol=1e-11, atol=1e-12)
        np.testing.assert_allclose(total, np.sum((want - y) ** 2), rtol=1e-11)
    # The scalar exchange: sums and maxima in rank order, identical on every rank, over three epochs.
    for epoch in (1, 2, 3):
        values = rng.standard_normal((world, 4))
        out = np.zeros((world, 4))
        for mx in (0, 1):
            assert lib.emu_peer_sums(world, ctypes.c_longlong(S), table, ctypes.c_ulonglong(2 * epoch - 1 + mx), 4, mx, p(values), p(out)) == 0
            ref = values[0].copy()
            for r in range(1, world):
                ref = np.maximum(ref, values[r]) if mx else ref + values[r]
            for r in range(world):
                np.testing.assert_array_equal(out[r], ref)


# ---- termination statistics (the two fused KKT passes) against the oracle ----

def _build_kkt():
    out = os.path.join(BUILD, "kkt_emu.so")
    sources = [os.path.join(EMU, "kkt_emu.cc"), os.path.join(EMU, "cuda_emu.h"), os.path.join(CSRC, "pdlp_skeletons.cuh"), os.path.join(CSRC, "pdlp_kkt_kernels.cuh")]
    os.makedirs(BUILD, exist_ok=True)
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(s) for s in sources):
        subprocess.run(["g++", "-std=c++20", "-O1", "-pthread", "-shared", "-fPIC", "-DPDLP_EMU_FIBERS", "-Wno-unknown-pragmas", "-Wno-attributes", "-I" + EMU, "-I" + CSRC,
                        sources[0], "-o", out], check=True)
    return ctypes.CDLL(out)


@pytest.mark.parametrize("qp_case, flag", [(False, True), (False, False), (True, True)])
def test_kkt_passes_match_the_oracles_convergence_and_infeasibility_information(qp_case, flag):
    """KktRowOp / KktColOp (pdlp_kkt_kernels.cuh) through the emulated row-stream kernel; the totals are
    assembled as PreprocessSolver::ComputeConvergenceAndInfeasibility does (pdlp_preprocess.cu) and
    compared with the oracle's ComputeConvergenceInformation / ComputeInfeasibilityInformation
    (iteration_stats.cc:349-518) on a point that sits on bounds, off bounds and at zero."""
    import scipy.sparse as sp
    import oracle_binding as ob
    K, c, q, vlb, vub, clb, cub, x, y = step_problem(11 + qp_case, qp_case)
    rng = np.random.default_rng(3)
    m, n = K.shape
    x = x + rng.standard_normal(n) * (rng.random(n) < 0.5)      # half the variables leave their projected start
    on = rng.random(n) < 0.2
    x[on & np.isfinite(vlb)] = vlb[on & np.isfinite(vlb)]       # some exactly on a bound
    x[rng.random(n) < 0.1] = 0.0
    y = rng.standard_normal(m) * (rng.random(m) < 0.7)          # zeros included
    row_scale, col_scale = rng.uniform(0.5, 2.0, m), rng.uniform(0.5, 2.0, n)
    x_start, y_start = np.minimum(np.maximum(np.zeros(n), vlb), vub), np.zeros(m)
    cw_p, cw_d = 2.5, 0.4

    qp = pdlp_b200.QuadraticProgram()
    qp.constraint_matrix = sp.csc_matrix(K)
    qp.objective_vector = c
    qp.variable_lower_bounds, qp.variable_upper_bounds = vlb, vub
    qp.constraint_lower_bounds, qp.constraint_upper_bounds = clb, cub
    if q is not None:
        qp.set_objective_matrix_diagonal(q)
    params = pdlp_b200.PrimalDualHybridGradientParams()
    params.handle_some_primal_gradients_on_finite_bounds_as_residuals = flag
    oracle = ob.Problem(qp)
    want_ci = oracle.convergence_info(params, col_scale, row_scale, x, y, cw_p, cw_d)
    want_ii = oracle.infeasibility_info(params, col_scale, row_scale, x, y)

    lib = _build_kkt()
    csr, csc = sp.csr_matrix(K), sp.csc_matrix(K)
    csr.sort_indices(); csc.sort_indices()
    i32 = lambda a: np.ascontiguousarray(a, dtype=np.int32)
    f64 = lambda a: np.ascontiguousarray(a, dtype=np.float64)
    arr = [i32(csr.indptr), i32(csr.indices), f64(csr.data)]
    rb = i32(pdlp_b200.work_blocks(arr[0])[0])
    arc = [i32(csc.indptr), i32(csc.indices), f64(csc.data)]
    cb = i32(pdlp_b200.work_blocks(arc[0])[0])
    p = lambda a: ctypes.c_void_p(a.ctypes.data) if a is not None else ctypes.c_void_p(None)
    vec = [f64(v) if v is not None else None for v in (c, q, vlb, vub, clb, cub, row_scale, col_scale, x, y, x_start, y_start)]
    r, t = np.zeros(9), np.zeros(15)
    assert lib.emu_kkt_totals(n, m, p(arr[0]), p(arr[1]), p(arr[2]), p(rb), len(rb) - 1, p(arc[0]), p(arc[1]), p(arc[2]), p(cb), len(cb) - 1, *[p(v) for v in vec],
                              ctypes.c_double(cw_p), ctypes.c_double(cw_d), int(flag), 3, p(r), p(t)) == 0
    nn = lambda v: max(v, 0.0)  # max slots start at 0 in the reference
    quad = 0.0 if q is None else 0.5 * t[1]
    dual_piece = -quad + r[3]
    got_ci = dict(l_inf_primal_residual=nn(r[1]), l2_primal_residual=np.sqrt(r[0]), l_inf_componentwise_primal_residual=nn(r[2]), l_inf_primal_variable=nn(t[7]),
                  l2_primal_variable=np.sqrt(t[8]), l_inf_dual_variable=nn(r[4]), l2_dual_variable=np.sqrt(r[5]), primal_objective=quad + t[0],
                  dual_objective=dual_piece + t[2], corrected_dual_objective=dual_piece + t[3], l_inf_dual_residual=nn(t[4]), l2_dual_residual=np.sqrt(t[5]),
                  l_inf_componentwise_dual_residual=nn(t[6]))
    for key, value in got_ci.items():
        np.testing.assert_allclose(value, want_ci.get(key, 0.0), rtol=1e-11, atol=1e-12, err_msg=key)
    linf_x, linf_y = nn(t[7]), nn(r[4])
    got_ii = dict(dual_ray_objective=(r[3] + t[9]) / linf_y, max_dual_ray_infeasibility=nn(t[10]) / linf_y, primal_ray_quadratic_norm=nn(t[12]) / linf_x,
                  max_primal_ray_infeasibility=max(nn(r[6]), nn(t[11])) / linf_x, primal_ray_linear_objective=t[0] / linf_x)
    for key, value in got_ii.items():
        np.testing.assert_allclose(value, want_ii.get(key, 0.0), rtol=1e-11, atol=1e-12, err_msg=key)
    # active-set counts of SetActiveSetInformation (primal_dual_hybrid_gradient.cc:1343-1412)
    assert t[13] == np.sum((x > vlb) & (x < vub))
    free = np.isneginf(clb) & np.isposinf(cub)
    assert r[7] == np.sum((y != 0) | free)

    # ReducedCosts (iteration_stats.cc:533-565)
    rc = np.zeros(n)
    for zero_objective in (0, 1):
        assert lib.emu_reduced_costs(n, p(arc[0]), p(arc[1]), p(arc[2]), p(cb), len(cb) - 1, p(vec[0]), p(vec[1]), p(vec[2]), p(vec[3]), p(vec[8]), p(vec[9]),
                                     zero_objective, int(flag), 3, p(rc)) == 0
        np.testing.assert_allclose(rc, oracle.reduced_costs(params, x, y, bool(zero_objective)), rtol=1e-11, atol=1e-12)

    # The split form used at host events (pdlp_kkt.cu: ComputeProducts + ComputeKktTotalsFromProducts): K x and K^T y of
    # THREE points with one stream of each matrix copy, then dense epilogue passes.  Every row sum is formed as the
    # single-point kernel forms it; the reduction slots add the same per-row terms in another order (a thread per row
    # instead of a lane per row group), so the totals agree to rounding and the max slots exactly.
    x_other, y_other = f64(rng.standard_normal(n)), f64(rng.standard_normal(m))
    r2, t2, kx, kty = np.zeros(9), np.zeros(15), np.zeros(3 * m), np.zeros(3 * n)
    assert lib.emu_kkt_totals_split(n, m, p(arr[0]), p(arr[1]), p(arr[2]), p(rb), len(rb) - 1, p(arc[0]), p(arc[1]), p(arc[2]), p(cb), len(cb) - 1, *[p(v) for v in vec],
                                    ctypes.c_double(cw_p), ctypes.c_double(cw_d), int(flag), 3, p(x_other), p(y_other), p(r2), p(t2), p(kx), p(kty)) == 0
    np.testing.assert_allclose(r2, r, rtol=1e-13, atol=0)
    np.testing.assert_allclose(t2, t, rtol=1e-13, atol=0)
    for s in (1, 2, 4, 6):
        assert r2[s] == r[s]  # max slots
    for s in (4, 6, 7, 10, 11, 12):
        assert t2[s] == t[s]
    for k, (xv, yv) in enumerate(((vec[8], vec[9]), (x_other, y_other), (vec[10], vec[11]))):
        np.testing.assert_allclose(kx[k * m:(k + 1) * m], csr @ xv, rtol=1e-12, atol=1e-12)
        np.testing.assert_allclose(kty[k * n:(k + 1) * n], csc.T @ yv, rtol=1e-12, atol=1e-12)


# ---- restart criteria: localized Lagrangian bounds against the oracle ----

def _build_restart():
    out = os.path.join(BUILD, "restart_emu.so")
    sources = [os.path.join(EMU, "restart_emu.cc"), os.path.join(EMU, "cuda_emu.h"), os.path.join(CSRC, "pdlp_skeletons.cuh"),
               os.path.join(CSRC, "pdlp_restart_kernels.cuh"), os.path.join(CSRC, "pdlp_restart_types.h")]
    os.makedirs(BUILD, exist_ok=True)
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(s) for s in sources):
        subprocess.run(["g++", "-std=c++20", "-O1", "-pthread", "-shared", "-fPIC", "-DPDLP_EMU_FIBERS", "-Wno-unknown-pragmas", "-Wno-attributes", "-I" + EMU, "-I" + CSRC,
                        sources[0], "-o", out], check=True)
    return ctypes.CDLL(out)


@pytest.mark.parametrize("qp_case, diagonal", [(False, False), (False, True), (True, True)])
def test_localized_lagrangian_bounds_match_the_oracle(qp_case, diagonal):
    """ComputeLocalizedLagrangianBounds, Euclidean norm (trust_region.cc:886-1016): the gradient ops,
    the trust-region ops and the root-finding host loop of pdlp_restart_kernels.cuh - the code the device
    path runs, driven through the emulated kernels - against the oracle's median-based solver, from
    radius 0 to a radius that fixes most variables at their bounds."""
    import scipy.sparse as sp
    import oracle_binding as ob
    K, c, q, vlb, vub, clb, cub, x, y = step_problem(21 + qp_case, qp_case)
    m, n = K.shape
    qp = pdlp_b200.QuadraticProgram()
    qp.constraint_matrix = sp.csc_matrix(K)
    qp.objective_vector = c
    qp.variable_lower_bounds, qp.variable_upper_bounds = vlb, vub
    qp.constraint_lower_bounds, qp.constraint_upper_bounds = clb, cub
    if q is not None:
        qp.set_objective_matrix_diagonal(q)
    oracle = ob.Problem(qp)
    lib = _build_restart()
    csr, csc = sp.csr_matrix(K), sp.csc_matrix(K)
    csr.sort_indices(); csc.sort_indices()
    i32 = lambda a: np.ascontiguousarray(a, dtype=np.int32)
    f64 = lambda a: np.ascontiguousarray(a, dtype=np.float64)
    arr = [i32(csr.indptr), i32(csr.indices), f64(csr.data)]
    rb = i32(pdlp_b200.work_blocks(arr[0])[0])
    arc = [i32(csc.indptr), i32(csc.indices), f64(csc.data)]
    cb = i32(pdlp_b200.work_blocks(arc[0])[0])
    p = lambda a: ctypes.c_void_p(a.ctypes.data) if a is not None else ctypes.c_void_p(None)
    vec = [f64(v) if v is not None else None for v in (c, q, vlb, vub, clb, cub, x, y)]
    tol = 1e-6 if diagonal else 1e-10  # (the diagonal solver bisects to `tol`: ~25 passes per radius at 1e-6)
    passes = []
    for radius, w in (((0.0, 1.0), (1.0, 0.6), (300.0, 1.0)) if diagonal else ((0.0, 1.0), (0.05, 1.0), (1.0, 0.6), (7.0, 2.0), (300.0, 1.0))):
        want = oracle.localized_bounds(x, y, 1, w, radius, diagonal, tol)
        out = np.zeros(5)
        assert lib.emu_localized_bounds(n, m, p(arr[0]), p(arr[1]), p(arr[2]), p(rb), len(rb) - 1, p(arc[0]), p(arc[1]), p(arc[2]), p(cb), len(cb) - 1,
                                        *[p(v) for v in vec], ctypes.c_double(w), ctypes.c_double(radius), int(diagonal), ctypes.c_double(tol), 3, p(out)) == 0
        got = dict(lagrangian_value=out[0], lower_bound=out[1], upper_bound=out[2], radius=out[3])
        scale = max(1.0, abs(want["lagrangian_value"]))
        for key in want:
            assert abs(got[key] - want[key]) <= (1e-5 if diagonal else 1e-10) * scale, (radius, key, got[key], want[key])
        passes.append(int(out[4]))
    assert passes[0] == 0 and max(passes) >= 2
    # The radius as the solver gets it: the weighted distance from the last restart point, reduced inside the gradient
    # passes (DistanceTraveledFromLastStart, primal_dual_hybrid_gradient.cc:1795-1804); the radius argument is ignored.
    rng = np.random.default_rng(99)
    xs, ys = f64(x + 0.3 * rng.standard_normal(n)), f64(y + 0.3 * rng.standard_normal(m))
    w = 0.6
    radius = float(np.sqrt(0.5 * w * np.sum((x - xs) ** 2) + 0.5 / w * np.sum((y - ys) ** 2)))
    want = oracle.localized_bounds(x, y, 1, w, radius, diagonal, tol)
    lib.emu_set_start_point(p(xs), p(ys))
    lib.emu_set_precomputed_kx(1)  # and K x handed in, as at a host event that computed the products of all its points at once
    try:
        out = np.zeros(5)
        assert lib.emu_localized_bounds(n, m, p(arr[0]), p(arr[1]), p(arr[2]), p(rb), len(rb) - 1, p(arc[0]), p(arc[1]), p(arc[2]), p(cb), len(cb) - 1,
                                        *[p(v) for v in vec], ctypes.c_double(w), ctypes.c_double(-1.0), int(diagonal), ctypes.c_double(tol), 3, p(out)) == 0
    finally:
        lib.emu_set_start_point(p(None), p(None))
        lib.emu_set_precomputed_kx(0)
    assert abs(out[3] - radius) <= 1e-12 * radius
    for key, got in zip(("lagrangian_value", "lower_bound", "upper_bound"), out[:3]):
        assert abs(got - want[key]) <= (1e-5 if diagonal else 1e-10) * max(1.0, abs(want["lagrangian_value"])), ("start point", key, got, want[key])
    if not diagonal:  # PrimalDualNorm::kMaxNorm: the primal and the dual problem solved separately (trust_region.cc:797-884)
        lib.emu_set_max_norm(1)
        try:
            for radius, w in ((0.0, 1.0), (0.05, 1.0), (1.0, 0.6), (7.0, 2.0), (300.0, 1.0)):
                want = oracle.localized_bounds(x, y, 0, w, radius, False, tol)
                out = np.zeros(5)
                assert lib.emu_localized_bounds(n, m, p(arr[0]), p(arr[1]), p(arr[2]), p(rb), len(rb) - 1, p(arc[0]), p(arc[1]), p(arc[2]), p(cb), len(cb) - 1,
                                                *[p(v) for v in vec], ctypes.c_double(w), ctypes.c_double(radius), 0, ctypes.c_double(tol), 3, p(out)) == 0
                got = dict(lagrangian_value=out[0], lower_bound=out[1], upper_bound=out[2], radius=out[3])
                scale = max(1.0, abs(want["lagrangian_value"]))
                for key in want:
                    assert abs(got[key] - want[key]) <= 1e-10 * scale, ("max norm", radius, key, got[key], want[key])
        finally:
            lib.emu_set_max_norm(0)


# ---- setup: Ruiz / L2 rescaling against the oracle ----

@pytest.mark.parametrize("qp_case, ruiz, l2", [(False, 5, True), (True, 3, False), (False, 0, True)])
def test_rescaling_matches_the_oracle(qp_case, ruiz, l2):
    """ApplyRescaling (sharded_optimization_utils.cc:367-444) and RescaleQuadraticProgram
    (sharded_quadratic_program.cc:97-167): scaled column / row norms over both matrix copies, the
    square-root updates, then the scaling of the vectors and of both copies of K."""
    import scipy.sparse as sp
    import oracle_binding as ob
    K, c, q, vlb, vub, clb, cub, _, _ = step_problem(31 + qp_case, qp_case)
    m, n = K.shape
    qp = pdlp_b200.QuadraticProgram()
    qp.constraint_matrix = sp.csc_matrix(K)
    qp.objective_vector = c
    qp.variable_lower_bounds, qp.variable_upper_bounds = vlb, vub
    qp.constraint_lower_bounds, qp.constraint_upper_bounds = clb, cub
    if q is not None:
        qp.set_objective_matrix_diagonal(q)
    oracle = ob.Problem(qp)
    want_row, want_col = oracle.rescale(ruiz, l2)
    want = oracle.get()

    out = os.path.join(BUILD, "setup_emu.so")
    sources = [os.path.join(EMU, "setup_emu.cc"), os.path.join(EMU, "cuda_emu.h"), os.path.join(CSRC, "pdlp_skeletons.cuh"), os.path.join(CSRC, "pdlp_ops_kernels.cuh")]
    os.makedirs(BUILD, exist_ok=True)
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(f) for f in sources):
        subprocess.run(["g++", "-std=c++20", "-O1", "-pthread", "-shared", "-fPIC", "-DPDLP_EMU_FIBERS", "-Wno-unknown-pragmas", "-Wno-attributes", "-I" + EMU, "-I" + CSRC,
                        sources[0], "-o", out], check=True)
    lib = ctypes.CDLL(out)
    csr, csc = sp.csr_matrix(K), sp.csc_matrix(K)
    csr.sort_indices(); csc.sort_indices()
    i32 = lambda a: np.ascontiguousarray(a, dtype=np.int32)
    f64 = lambda a: np.array(a, dtype=np.float64)
    rp, ri, rv = i32(csr.indptr), i32(csr.indices), f64(csr.data)
    cp, ci, cv = i32(csc.indptr), i32(csc.indices), f64(csc.data)
    rb, cb = i32(pdlp_b200.work_blocks(rp)[0]), i32(pdlp_b200.work_blocks(cp)[0])
    vec = dict(c=f64(c), q=None if q is None else f64(q), vlb=f64(vlb), vub=f64(vub), clb=f64(clb), cub=f64(cub))
    row, col = np.ones(m), np.ones(n)
    p = lambda a: ctypes.c_void_p(a.ctypes.data) if a is not None else ctypes.c_void_p(None)
    assert lib.emu_rescale(n, m, p(rp), p(ri), p(rv), p(rb), len(rb) - 1, p(cp), p(ci), p(cv), p(cb), len(cb) - 1, p(vec["c"]), p(vec["q"]), p(vec["vlb"]),
                           p(vec["vub"]), p(vec["clb"]), p(vec["cub"]), ruiz, int(l2), 3, p(row), p(col)) == 0
    np.testing.assert_allclose(row, want_row, rtol=1e-12)
    np.testing.assert_allclose(col, want_col, rtol=1e-12)
    for key in ("c", "vlb", "vub", "clb", "cub") + (("q",) if q is not None else ()):
        np.testing.assert_allclose(vec[key], want[key], rtol=1e-12, err_msg=key)
    np.testing.assert_allclose(cv, want["val"], rtol=1e-12)  # the oracle's matrix is the CSC copy
    np.testing.assert_allclose(sp.csr_matrix((rv, ri, rp), shape=(m, n)).toarray(), sp.csc_matrix((cv, ci, cp), shape=(m, n)).toarray(), rtol=1e-13)


# ---- shared-memory protocol under a race detector ----

def test_rowstream_shared_memory_protocol_is_race_free():
    """tests/kernel_emu/race_check.cc under ThreadSanitizer: with __syncthreads as the only
    cross-thread ordering, every conflicting pair of accesses to the staged products, the row offsets
    and the single index buffer (this block's indices read, the next block's copied into the same
    positions) must be separated by a barrier, in both cp.async orders.  Removing one barrier from the
    kernel makes this fail (checked by hand when the test was written)."""
    exe = os.path.join(BUILD, "race_check")
    src = os.path.join(EMU, "race_check.cc")
    os.makedirs(BUILD, exist_ok=True)
    build = subprocess.run(["g++", "-std=c++20", "-O1", "-g", "-fsanitize=thread", "-pthread", "-Wno-unknown-pragmas", "-Wno-attributes",
                            "-I" + EMU, "-I" + CSRC, src, "-o", exe], capture_output=True, text=True)
    if build.returncode != 0 and ("tsan" in build.stderr.lower() or "sanitize" in build.stderr.lower()):
        pytest.skip("ThreadSanitizer runtime not available: " + build.stderr.strip().splitlines()[-1])
    assert build.returncode == 0, build.stderr
    run = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    text = run.stdout + run.stderr
    if "FATAL: ThreadSanitizer" in text:
        pytest.skip("ThreadSanitizer cannot run here: " + text.strip().splitlines()[0])
    assert "ThreadSanitizer" not in text, text[:4000]
    assert run.returncode == 0 and "wrong rows: 0" in text, text[:2000]


@pytest.mark.parametrize("shape", ["no_rows", "no_columns", "empty_rows_and_columns"])
def test_adaptive_trial_steps_on_degenerate_shapes(step_libs, shape):
    """A problem with variables and no constraint (model_builder_helper_test.py:104-126 solves one through PDLP), its
    mirror, and a matrix whose only nonzero leaves whole rows and columns empty: the kernels must run over zero work
    blocks / zero-length rows and the step rule must see nonlinearity 0 (limit = inf: accept, grow)."""
    import scipy.sparse as sp
    if shape == "no_rows":
        K = sp.csr_matrix((0, 3))
    elif shape == "no_columns":
        K = sp.csr_matrix((2, 0))
    else:
        K = sp.csr_matrix(([2.0], ([1], [2])), shape=(4, 5))
    m, n = K.shape
    rng = np.random.default_rng(5)
    c = rng.standard_normal(n)
    vlb, vub = np.full(n, -1.0), np.full(n, 2.0)
    # Rows that 0 = (K x)_i cannot satisfy keep y moving (a feasible row pulls y to 0 and the movement to 0 with it).
    clb, cub = np.full(m, 0.5), np.full(m, 0.5)
    x0, y0 = np.zeros(n), rng.standard_normal(m)
    step, w, trials = 0.3, 1.5, 6
    want = reference_adaptive_trials(K, c, None, vlb, vub, clb, cub, x0, y0, step, w, trials)
    got = emulated_adaptive_trials(step_libs[0], K, c, None, vlb, vub, clb, cub, x0, y0, step, w, trials)
    # want["rejected"] counts a rejected trial still pending at the end of the batch; the device adds those to
    # num_rejected_steps at the next accepted step (pdlp_step_state.h: inner), so compare through the trial total.
    assert got["active"] == 1 and got["total_trials"] == trials
    assert (got["completed"], got["total_trials"] - got["completed"]) == (want["completed"], want["rejected"])
    for key in ("x", "y", "avg_x", "avg_y"):
        np.testing.assert_allclose(got[key], want[key], rtol=1e-12, atol=1e-13, err_msg=key)
    for key in ("step", "movement", "nonlinearity", "weight_sum"):
        np.testing.assert_allclose(got[key], want[key], rtol=1e-12, err_msg=key)

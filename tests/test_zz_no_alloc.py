"""pdlp_b200_matvec_no_alloc (include/pdlp_b200.h): K v from caller-owned DEVICE arrays with the work-block
table in a static device array - no cudaMalloc inside the call.  No counterpart in the reference,
so the check is against the existing entry point,
pdlp_b200_matvec(which = 1), on the same matrix, to the 1e-12 relative tolerance the kernel-level tests use
(the two tables are cut differently, so rows shared by several lanes are summed in another lane order).

Two calls in a row on different matrices (the second must not see anything of the first in the shared
scratch array), then the first matrix again.  Collected last on purpose:
`pytest --collect-only -q tests` ends with this file.  Device memory for the test comes from torch; the
entry point itself allocates nothing."""
import ctypes

import numpy as np
import pytest
import scipy.sparse as sp

import pdlp_b200

pytestmark = pytest.mark.gpu


def csr_case(rows, cols, lengths, seed):
    rng = np.random.default_rng(seed)
    lengths = np.asarray(lengths, dtype=np.int64)
    ptr = np.concatenate([[0], np.cumsum(lengths)]).astype(np.int32)
    idx = np.concatenate([np.sort(rng.choice(cols, size=int(k), replace=False)) for k in lengths] + [np.zeros(0, dtype=np.int64)]).astype(np.int32)
    val = rng.standard_normal(int(ptr[-1]))
    return sp.csr_matrix((val, idx, ptr), shape=(rows, cols)), rng.standard_normal(cols)


def existing_entry_point(K, v):
    qp = pdlp_b200.QuadraticProgram()
    qp.resize_and_initialize(K.shape[1], K.shape[0])
    qp.constraint_matrix = sp.csc_matrix(K)
    dp = pdlp_b200.DeviceProblem(qp)
    try:
        return dp.matvec(1, v)[0]
    finally:
        dp.close()


def no_alloc(K, v, torch):
    dev = torch.device("cuda", 0)
    ptr = torch.from_numpy(np.ascontiguousarray(K.indptr, dtype=np.int32)).to(dev)
    idx = torch.from_numpy(np.ascontiguousarray(K.indices, dtype=np.int32)).to(dev)
    val = torch.from_numpy(np.ascontiguousarray(K.data, dtype=np.float64)).to(dev)
    vd = torch.from_numpy(np.ascontiguousarray(v, dtype=np.float64)).to(dev)
    out = torch.full((K.shape[0],), float("nan"), dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(0)[0]  # the driver's count: sees any cudaMalloc, torch's or not
    status = pdlp_b200._lib.pdlp_b200_matvec_no_alloc(K.shape[0], K.shape[1], ptr.data_ptr(), idx.data_ptr(), val.data_ptr(), vd.data_ptr(), out.data_ptr(), 0, None)
    assert status == 0, pdlp_b200._lib.pdlp_b200_last_error().decode()
    free_after = torch.cuda.mem_get_info(0)[0]
    return out.cpu().numpy(), free_before - free_after


def test_no_alloc_matvec_matches_the_existing_entry_point_on_two_calls_in_a_row():
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(5)
    # A: short uniform rows (1024 rows per block fit).  B: ragged rows with a few longer than a work block
    # (forces few rows per block and the long-row path), fewer rows than A so A's table tail stays behind it.
    a = csr_case(30000, 5000, np.full(30000, 2), 1)
    lengths = rng.poisson(40, 9000)
    lengths[[17, 4000, 8999]] = [2500, 2048, 3000]
    b = csr_case(9000, 4000, lengths, 2)
    for call, (K, v) in enumerate((a, b, a)):
        want = existing_entry_point(K, v)
        got, consumed = no_alloc(K, v, torch)
        # The first call may load the module that holds the static array (lazy loading); later calls must
        # leave the device's free memory exactly where it was.
        assert call == 0 or consumed == 0, consumed
        scale = np.abs(K) @ np.abs(v) + 1e-300
        assert np.max(np.abs(got - want) / scale) <= 1e-12
        np.testing.assert_allclose(got, K @ v, rtol=0, atol=1e-12 * float(scale.max()))


def test_no_alloc_matvec_rejects_sizes_beyond_the_static_scratch():
    rows = pdlp_b200.NO_ALLOC_MAX_ROWS + 1
    dummy = ctypes.c_void_p(8)  # never dereferenced: the size check comes first
    status = pdlp_b200._lib.pdlp_b200_matvec_no_alloc(rows, 4, dummy, dummy, dummy, dummy, dummy, 0, None)
    assert status == 1
    msg = pdlp_b200._lib.pdlp_b200_last_error().decode()
    assert str(pdlp_b200.NO_ALLOC_MAX_ROWS) in msg and "exceed" in msg

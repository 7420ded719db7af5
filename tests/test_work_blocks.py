"""Host-side invariants of the work-block partition the row-stream kernels run on
(pdlp_matrix.cu:ComputeBlocks through pdlp_b200_work_blocks).  The kernels stage
one block in shared memory, so a block that broke these limits would corrupt a solve rather than
fail: at most 2048 nonzeros unless it is one long row, at most 1024 rows, no block across the
shared / private cut of a sharded constraint block."""
import numpy as np
import pytest

import pdlp_b200

CHUNK, MAX_ROWS = 2048, 1024


def offsets_of(lengths):
    return np.concatenate([[0], np.cumsum(np.asarray(lengths, dtype=np.int64))])


def check_partition(lengths, starts, first=0, last=None):
    lengths = np.asarray(lengths, dtype=np.int64)
    last = len(lengths) if last is None else last
    assert starts[0] == first and starts[-1] == last
    assert np.all(np.diff(starts) > 0) or last == first
    for b in range(len(starts) - 1):
        r0, r1 = starts[b], starts[b + 1]
        nz = lengths[r0:r1].sum()
        assert r1 - r0 <= MAX_ROWS
        assert nz <= CHUNK or r1 - r0 == 1, (b, r0, r1, nz)
        # greedy: the next row did not fit (or the block was full of rows, or a long row stands alone)
        if r1 < last:
            assert nz > CHUNK or r1 - r0 == MAX_ROWS or nz + lengths[r1] > CHUNK, (b, r0, r1)


SHAPES = {
    "uniform16": lambda rng: np.full(5000, 16),
    "uniform3": lambda rng: np.full(7001, 3),
    "poisson10": lambda rng: rng.poisson(10, 4000),
    "empty_rows": lambda rng: rng.integers(0, 2, 6000),
    "all_empty": lambda rng: np.zeros(3000, dtype=np.int64),
    "long_rows": lambda rng: np.where(rng.random(300) < 0.1, rng.integers(2049, 9000, 300), rng.integers(0, 40, 300)),
    "exactly_chunk": lambda rng: np.array([2048, 2048, 1, 2047, 1, 2049, 0, 0, 5]),
    "one_row": lambda rng: np.array([7]),
}


@pytest.mark.parametrize("shape", sorted(SHAPES))
def test_blocks_respect_the_staging_limits(shape):
    lengths = SHAPES[shape](np.random.default_rng(3))
    starts, cut_block, mx = pdlp_b200.work_blocks(offsets_of(lengths))
    check_partition(lengths, starts)
    assert cut_block == -1
    assert mx == lengths.max()


def test_no_rows_gives_no_blocks():
    starts, cut_block, mx = pdlp_b200.work_blocks([0])
    assert starts.tolist() == [0] and cut_block == -1 and mx == 0


@pytest.mark.parametrize("shape", ["uniform16", "poisson10", "empty_rows", "long_rows", "all_empty"])
def test_a_cut_is_a_block_boundary_and_both_sides_are_partitioned_independently(shape):
    rng = np.random.default_rng(11)
    lengths = SHAPES[shape](rng)
    off = offsets_of(lengths)
    rows = len(lengths)
    for cut in sorted({1, 2, rows // 3, rows // 2 + 1, rows - 1}):
        if not 0 < cut < rows:
            continue
        starts, cut_block, mx = pdlp_b200.work_blocks(off, cut)
        assert starts[cut_block] == cut
        assert mx == lengths.max()
        check_partition(lengths, starts[:cut_block + 1], 0, cut)
        check_partition(lengths, starts[cut_block:], cut, rows)
        head, _, _ = pdlp_b200.work_blocks(off[:cut + 1])
        tail, _, _ = pdlp_b200.work_blocks(off[cut:] - off[cut])
        np.testing.assert_array_equal(starts[:cut_block + 1], head)
        np.testing.assert_array_equal(starts[cut_block:], tail + cut)


def test_offsets_need_not_start_at_zero():
    lengths = np.random.default_rng(5).poisson(12, 3000)
    off = offsets_of(lengths)
    a, _, _ = pdlp_b200.work_blocks(off)
    b, _, _ = pdlp_b200.work_blocks(off + 12345)
    np.testing.assert_array_equal(a, b)


@pytest.mark.parametrize("offsets, cut", [([0, 5, 3], -1), ([0, 1, 2], 0), ([0, 1, 2], 2), ([0, 1, 2], 7)])
def test_bad_arguments_are_rejected(offsets, cut):
    with pytest.raises(ValueError):
        pdlp_b200.work_blocks(offsets, cut)

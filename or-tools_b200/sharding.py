"""Column sharding of a QuadraticProgram over GPUs, one process per GPU.

The reference shards every pass over the constraint matrix by contiguous column ranges of
balanced mass `1 + nnz(column)` across CPU threads (ShardedQuadraticProgram,
/root/reference/ortools/pdlp/sharded_quadratic_program.h:38-60; Sharder's mass constructor,
sharder.cc:36-70, and the matrix overload sharder.h:161-165).  The same ranges, one per
rank, are the unit of multi-GPU work here: rank r keeps columns [starts[r], starts[r+1]) of K,
the matching slices of the primal-side vectors, and all dual-side vectors.

Host-side only (numpy / scipy); the numeric work is pdlp_b200_solve_sharded in the C ABI.
"""
import ctypes

import numpy as np

from . import UNIQUE_ID_BYTES, _check, _CShard, _lib
from ._api import QuadraticProgram, primal_dual_hybrid_gradient

try:
    import scipy.sparse as _sp
except ImportError:  # pragma: no cover
    _sp = None


def column_partition(col_ptr, world):
    """Boundaries (world + 1 int64, first 0, last n) of `world` non-empty contiguous column blocks
    whose masses sum(1 + nnz(col)) are balanced.

    The element mass is the reference's (sharder.h:161-165).  Its greedy boundary rule
    (sharder.cc:51-63) may return fewer or more shards than asked, which a fixed set of ranks
    cannot accept, so the cut here is on the prefix mass: block k ends at the first column where
    the prefix reaches (k + 1) / world of the total, then boundaries are nudged so that no block
    is empty.  Requires n >= world."""
    col_ptr = np.asarray(col_ptr, dtype=np.int64)
    n = col_ptr.size - 1
    if world < 1:
        raise ValueError("world must be >= 1")
    if n < world:
        raise ValueError(f"cannot give each of {world} ranks a column: the problem has {n}")
    prefix = col_ptr[1:] - col_ptr[0] + np.arange(1, n + 1, dtype=np.int64)  # mass of columns [0, j]
    total = int(prefix[-1])
    starts = np.zeros(world + 1, dtype=np.int64)
    starts[world] = n
    for k in range(1, world):
        target = -(-total * k // world)  # ceil
        starts[k] = int(np.searchsorted(prefix, target, side="left")) + 1
    for k in range(1, world):  # non-empty, increasing
        starts[k] = min(max(starts[k], starts[k - 1] + 1), n - (world - k))
    return starts


def take_columns(qp, lo, hi):
    """The column block [lo, hi) of `qp` as a QuadraticProgram: those columns of K (all m rows),
    the slices of objective vector / diagonal / variable bounds, and the whole constraint bounds.
    The objective offset stays whole on every rank (it is added once, after the reductions)."""
    K = qp.constraint_matrix
    if _sp is not None and not _sp.isspmatrix_csc(K):
        K = _sp.csc_matrix(K)
    out = QuadraticProgram()
    out.constraint_matrix = K[:, lo:hi]
    out.objective_vector = np.asarray(qp.objective_vector, dtype=np.float64)[lo:hi].copy()
    out.variable_lower_bounds = np.asarray(qp.variable_lower_bounds, dtype=np.float64)[lo:hi].copy()
    out.variable_upper_bounds = np.asarray(qp.variable_upper_bounds, dtype=np.float64)[lo:hi].copy()
    out.constraint_lower_bounds = np.asarray(qp.constraint_lower_bounds, dtype=np.float64).copy()
    out.constraint_upper_bounds = np.asarray(qp.constraint_upper_bounds, dtype=np.float64).copy()
    if qp.objective_matrix is not None:
        out.set_objective_matrix_diagonal(np.asarray(qp.objective_matrix, dtype=np.float64)[lo:hi].copy())
    out.objective_offset = qp.objective_offset
    out.objective_scaling_factor = qp.objective_scaling_factor
    out.problem_name = qp.problem_name
    return out


def unique_id():
    """A fresh NCCL unique id (bytes); call on ONE rank and hand it to the others."""
    buf = (ctypes.c_uint8 * UNIQUE_ID_BYTES)()
    _check(_lib.pdlp_b200_unique_id(buf))
    return bytes(buf)


class ColumnShard:
    """What pdlp_b200_shard carries: which block of the global columns this rank holds."""

    def __init__(self, rank, world, global_num_variables, column_offset, nccl_unique_id):
        if world > 1 and (nccl_unique_id is None or len(nccl_unique_id) != UNIQUE_ID_BYTES):
            raise ValueError(f"nccl_unique_id must be {UNIQUE_ID_BYTES} bytes")
        self.rank, self.world = int(rank), int(world)
        self.global_num_variables, self.column_offset = int(global_num_variables), int(column_offset)
        self._id = (ctypes.c_uint8 * UNIQUE_ID_BYTES).from_buffer_copy(nccl_unique_id) if nccl_unique_id is not None else None
        self.c_shard = _CShard(self.rank, self.world, self.global_num_variables, self.column_offset,
                               ctypes.cast(self._id, ctypes.c_void_p) if self._id is not None else None)


def broadcast_unique_id(group=None):
    """Rank 0 of the torch.distributed group draws the id; everybody returns the same bytes."""
    import torch
    import torch.distributed as dist

    rank = dist.get_rank(group)
    payload = [unique_id() if rank == 0 else None]
    dist.broadcast_object_list(payload, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    del torch
    return payload[0]


def primal_dual_hybrid_gradient_sharded(qp, params, initial_solution=None, interrupt_solve=None, iteration_stats_callback=None, device=-1,
                                        group=None, starts=None, gather=True):
    """PrimalDualHybridGradient over the ranks of a torch.distributed group, one GPU each.

    Every rank passes the same whole `qp` (or `starts` plus a qp that is already its own block --
    then `qp.global_num_variables` is not needed: starts[-1] is the global count) and receives the
    same SolverResult: with gather=True the primal solution and reduced costs are all-gathered to
    full length, otherwise they are this rank's block."""
    import torch.distributed as dist

    rank, world = dist.get_rank(group), dist.get_world_size(group)
    if world == 1:
        return primal_dual_hybrid_gradient(qp, params, initial_solution, interrupt_solve, iteration_stats_callback, device)
    K = qp.constraint_matrix
    whole = starts is None
    if whole:
        if _sp is not None and not _sp.isspmatrix_csc(K):
            K = _sp.csc_matrix(K)
        starts = column_partition(K.indptr, world)
        block = take_columns(qp, int(starts[rank]), int(starts[rank + 1]))
    else:
        block = qp
    lo, hi = int(starts[rank]), int(starts[rank + 1])
    if initial_solution is not None and whole:
        from ._api import PrimalAndDualSolution

        initial_solution = PrimalAndDualSolution(np.asarray(initial_solution.primal_solution)[lo:hi], initial_solution.dual_solution)
    shard = ColumnShard(rank, world, int(starts[-1]), lo, broadcast_unique_id(group))
    result = primal_dual_hybrid_gradient(block, params, initial_solution, interrupt_solve, iteration_stats_callback, device, shard=shard)
    if gather and result.primal_solution.size:
        parts = [None] * world
        dist.all_gather_object(parts, (result.primal_solution, result.reduced_costs), group=group)
        result.primal_solution = np.concatenate([p[0] for p in parts])
        result.reduced_costs = np.concatenate([p[1] for p in parts])
    return result

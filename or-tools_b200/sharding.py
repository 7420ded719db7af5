"""Column sharding of a QuadraticProgram over GPUs, one process per GPU.

The reference shards every pass over the constraint matrix by contiguous column ranges of
balanced mass `1 + nnz(column)` across CPU threads (ShardedQuadraticProgram,
/root/reference/ortools/pdlp/sharded_quadratic_program.h:38-60; Sharder's mass constructor,
sharder.cc:36-70, and the matrix overload sharder.h:161-165).  The same ranges, one per
rank, are the unit of multi-GPU work here: rank r keeps columns [starts[r], starts[r+1]) of K and
the matching slices of the primal-side vectors.  Of the dual side it keeps only the constraints
its columns touch: those touched by several ranks ("shared", exchanged by all-reduce every step)
first, then those touched by this rank alone ("private", no exchange at all).

Host-side only (numpy / scipy); the numeric work is pdlp_b200_solve_sharded in the C ABI.
"""
import ctypes

import numpy as np

from . import UNIQUE_ID_BYTES, _check, _CShard, _lib
from ._api import PrimalAndDualSolution, QuadraticProgram, primal_dual_hybrid_gradient

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


class RowSharing:
    """Which constraints each rank keeps, from which ranks' columns touch them.

    shared        global ids (ascending) of constraints touched by >= 2 ranks: on every rank, first
    private[r]    global ids (ascending) of constraints touched by rank r only; constraints no rank
                  touches (empty rows of K) go to rank 0
    local_rows(r) = concat(shared, private[r]): the rank's constraints in its local order
    """

    def __init__(self, shared, private, num_constraints):
        self.shared = np.asarray(shared, dtype=np.int64)
        self.private = [np.asarray(p, dtype=np.int64) for p in private]
        self.num_constraints = int(num_constraints)

    @classmethod
    def from_touch_counts(cls, touched_by_rank, num_constraints):
        """touched_by_rank[r]: bool array (m) or index array of the constraints rank r's columns touch."""
        world = len(touched_by_rank)
        count = np.zeros(num_constraints, dtype=np.int32)
        owner = np.zeros(num_constraints, dtype=np.int32)  # rank 0 owns what nobody touches
        for r, t in enumerate(touched_by_rank):
            t = np.asarray(t)
            idx = np.flatnonzero(t) if t.dtype == np.bool_ else t.astype(np.int64)
            count[idx] += 1
            owner[idx] = r
        shared = np.flatnonzero(count >= 2)
        single = count <= 1
        private = [np.flatnonzero(single & (owner == r)) for r in range(world)]
        return cls(shared, private, num_constraints)

    @classmethod
    def from_matrix(cls, K, starts):
        """K: the whole constraint matrix (CSC); starts: column_partition boundaries."""
        m = K.shape[0]
        touched = []
        for r in range(len(starts) - 1):
            rows = K.indices[K.indptr[starts[r]]:K.indptr[starts[r + 1]]]
            t = np.zeros(m, dtype=np.bool_)
            t[rows] = True
            touched.append(t)
        return cls.from_touch_counts(touched, m)

    @classmethod
    def replicated(cls, num_constraints, world):
        """Every rank keeps every constraint (all shared): the plain layout."""
        return cls(np.arange(num_constraints, dtype=np.int64), [np.zeros(0, dtype=np.int64)] * world, num_constraints)

    def local_rows(self, rank):
        return np.concatenate([self.shared, self.private[rank]])

    def scatter(self, rank, dual_vector):
        """Whole (m) dual-side vector -> this rank's local order."""
        return np.asarray(dual_vector, dtype=np.float64)[self.local_rows(rank)]

    def gather(self, local_vectors):
        """Per-rank local dual-side vectors -> whole (m) vector (the shared part from rank 0)."""
        out = np.zeros(self.num_constraints)
        out[self.shared] = local_vectors[0][: self.shared.size]
        for r, v in enumerate(local_vectors):
            out[self.private[r]] = v[self.shared.size:]
        return out


def take_columns(qp, lo, hi, rows=None):
    """The column block [lo, hi) of `qp` as a QuadraticProgram: those columns of K, the slices of
    objective vector / diagonal / variable bounds.  `rows` (global constraint ids in the rank's
    local order, RowSharing.local_rows) restricts and renumbers the constraints; None keeps all.
    The objective offset stays whole on every rank (it is added once, after the reductions)."""
    K = qp.constraint_matrix
    if _sp is not None and not _sp.isspmatrix_csc(K):
        K = _sp.csc_matrix(K)
    block = K[:, lo:hi]
    clb = np.asarray(qp.constraint_lower_bounds, dtype=np.float64)
    cub = np.asarray(qp.constraint_upper_bounds, dtype=np.float64)
    if rows is not None:
        rows = np.asarray(rows, dtype=np.int64)
        block = _sp.csc_matrix(block.tocsr()[rows, :])  # every nonzero of the block lies in `rows`
        clb, cub = clb[rows], cub[rows]
    out = QuadraticProgram()
    out.constraint_matrix = block
    out.objective_vector = np.asarray(qp.objective_vector, dtype=np.float64)[lo:hi].copy()
    out.variable_lower_bounds = np.asarray(qp.variable_lower_bounds, dtype=np.float64)[lo:hi].copy()
    out.variable_upper_bounds = np.asarray(qp.variable_upper_bounds, dtype=np.float64)[lo:hi].copy()
    out.constraint_lower_bounds = clb.copy()
    out.constraint_upper_bounds = cub.copy()
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
    """What pdlp_b200_shard carries: which block of the global columns this rank holds and how many
    of its leading constraints are shared with other ranks (None: all of them)."""

    def __init__(self, rank, world, global_num_variables, column_offset, nccl_unique_id, num_shared_constraints=None):
        if world > 1 and (nccl_unique_id is None or len(nccl_unique_id) != UNIQUE_ID_BYTES):
            raise ValueError(f"nccl_unique_id must be {UNIQUE_ID_BYTES} bytes")
        self.rank, self.world = int(rank), int(world)
        self.global_num_variables, self.column_offset = int(global_num_variables), int(column_offset)
        self.num_shared_constraints = -1 if num_shared_constraints is None else int(num_shared_constraints)
        self._id = (ctypes.c_uint8 * UNIQUE_ID_BYTES).from_buffer_copy(nccl_unique_id) if nccl_unique_id is not None else None
        self.c_shard = _CShard(self.rank, self.world, self.global_num_variables, self.column_offset,
                               ctypes.cast(self._id, ctypes.c_void_p) if self._id is not None else None, self.num_shared_constraints)


_group_ids = {}


def broadcast_unique_id(group=None, fresh=False):
    """Rank 0 of the torch.distributed group draws the id; everybody returns the same bytes.  The id
    is drawn once per group and process (the library keeps the communicator behind it, so later
    solves skip NCCL's set-up); fresh=True draws a new one (all ranks must pass the same flag)."""
    import torch.distributed as dist

    key = id(group) if group is not None else None
    if not fresh and key in _group_ids:
        return _group_ids[key]
    rank = dist.get_rank(group)
    payload = [unique_id() if rank == 0 else None]
    dist.broadcast_object_list(payload, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    _group_ids[key] = payload[0]
    return payload[0]


def release_communicators():
    """Destroys the NCCL communicators the library kept (collective; e.g. before the process group goes away)."""
    released = ctypes.c_int(0)
    _check(_lib.pdlp_b200_release_communicators(ctypes.byref(released)))
    _group_ids.clear()
    return released.value


def primal_dual_hybrid_gradient_sharded(qp, params, initial_solution=None, interrupt_solve=None, iteration_stats_callback=None, device=-1,
                                        group=None, starts=None, sharing=None, gather=True, private_rows=True, num_shared_constraints=None, unique_id=None):
    """PrimalDualHybridGradient over the ranks of a torch.distributed group, one GPU each.

    Whole-problem mode (starts is None): every rank passes the same whole `qp` and, with gather=True,
    receives the same whole SolverResult.  Block mode: `qp` is already this rank's block in its local
    row order, `starts` the column boundaries and `sharing` the RowSharing that produced the local
    order (None: every rank holds all constraints, or give num_shared_constraints when only the count
    is known); results stay local unless gather=True (which needs `sharing` for the dual vector).
    private_rows=False keeps every constraint on every rank (one all-reduce of all m row sums)."""
    import torch.distributed as dist

    rank, world = dist.get_rank(group), dist.get_world_size(group)
    if world == 1:
        return primal_dual_hybrid_gradient(qp, params, initial_solution, interrupt_solve, iteration_stats_callback, device)
    whole = starts is None
    if whole:
        K = qp.constraint_matrix
        if _sp is not None and not _sp.isspmatrix_csc(K):
            K = _sp.csc_matrix(K)
        starts = column_partition(K.indptr, world)
        sharing = RowSharing.from_matrix(K, starts) if private_rows else RowSharing.replicated(K.shape[0], world)
        lo, hi = int(starts[rank]), int(starts[rank + 1])
        block = take_columns(qp, lo, hi, sharing.local_rows(rank))
        if initial_solution is not None:
            initial_solution = PrimalAndDualSolution(np.asarray(initial_solution.primal_solution, dtype=np.float64)[lo:hi],
                                                     sharing.scatter(rank, initial_solution.dual_solution))
    else:
        block = qp
        lo = int(starts[rank])
    num_shared = num_shared_constraints if sharing is None else int(sharing.shared.size)
    # unique_id: the id of an earlier sharded call of this process group; its communicator (and its peer-memory
    # mappings) are cached per process and reused, so only the first call pays for their creation.
    shard = ColumnShard(rank, world, int(starts[-1]), lo, unique_id if unique_id is not None else broadcast_unique_id(group), num_shared)
    result = primal_dual_hybrid_gradient(block, params, initial_solution, interrupt_solve, iteration_stats_callback, device, shard=shard)
    if gather and result.primal_solution.size:
        parts = [None] * world
        dist.all_gather_object(parts, (result.primal_solution, result.reduced_costs, result.dual_solution), group=group)
        result.primal_solution = np.concatenate([p[0] for p in parts])
        result.reduced_costs = np.concatenate([p[1] for p in parts])
        if sharing is not None:
            result.dual_solution = sharing.gather([p[2] for p in parts])
    return result

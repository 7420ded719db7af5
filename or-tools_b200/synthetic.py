"""Synthetic problem generators for the benchmark configurations of BASELINE.json
(SURVEY.md 8d).  Host-side numpy; all take a seed and return a QuadraticProgram whose
optimum exists by construction (a primal-dual optimal pair is planted for the random
LP / QP; the network problems are feasible and bounded by positive costs)."""
import numpy as np
import scipy.sparse as sp

from ._qp import QuadraticProgram

INF = np.inf


def _csc_from_columns(m, n, rows, vals, per_col):
    """rows/vals: (n, per_col) arrays, rows sorted inside each column."""
    # 32-bit indices where they fit (scipy's own choice; the 1e9-nnz LP then needs 6 GB of index arrays instead of 12).
    idx = np.int32 if max(m, n * per_col) < 2 ** 31 else np.int64
    indptr = np.arange(0, n * per_col + 1, per_col, dtype=idx)
    return sp.csc_matrix((vals.reshape(-1), rows.reshape(-1).astype(idx, copy=False), indptr), shape=(m, n))


def random_sparse_lp(m, n, nnz_per_col=5, seed=0, equality_fraction=0.5, qp_diagonal=False):
    """Config C2 (m = 1e6, n = 2e6, 5 per column) and, with qp_diagonal, C4.

    K has exactly nnz_per_col distinct uniform rows per column, N(0,1) values; variables
    live in [0, U(1,10)]; half the rows are equalities, the rest `<=` rows.  A planted
    pair (x*, y*) satisfies the KKT conditions, so the optimum is known."""
    rng = np.random.Generator(np.random.PCG64(seed))
    k = min(nnz_per_col, m)
    rows = rng.integers(0, m, size=(n, k), dtype=np.int64)
    rows.sort(axis=1)
    for _ in range(64):  # resample duplicates inside a column
        dup = np.zeros_like(rows, dtype=bool)
        dup[:, 1:] = rows[:, 1:] == rows[:, :-1]
        bad = dup.any(axis=1)
        if not bad.any():
            break
        rows[bad] = rng.integers(0, m, size=(int(bad.sum()), k), dtype=np.int64)
        rows[bad] = np.sort(rows[bad], axis=1)
    vals = rng.standard_normal((n, k))
    K = _csc_from_columns(m, n, rows, vals, k)

    ub = rng.uniform(1.0, 10.0, n)
    lb = np.zeros(n)
    where = rng.integers(0, 3, n)  # 0: at lower, 1: at upper, 2: interior
    x_star = np.where(where == 0, lb, np.where(where == 1, ub, rng.uniform(0.0, 1.0, n) * ub))
    activity = K @ x_star
    is_eq = rng.uniform(size=m) < equality_fraction
    active_ineq = (~is_eq) & (rng.uniform(size=m) < 0.5)
    slack = np.where(is_eq | active_ineq, 0.0, rng.uniform(0.0, 1.0, m))
    cub = activity + slack
    clb = np.where(is_eq, activity, -INF)
    # Dual: free sign on equalities, <= 0 on active `<=` rows, 0 on slack rows.
    y_star = np.where(is_eq, rng.standard_normal(m), np.where(active_ineq, -np.abs(rng.standard_normal(m)), 0.0))
    reduced = np.where(where == 0, np.abs(rng.standard_normal(n)), np.where(where == 1, -np.abs(rng.standard_normal(n)), 0.0))
    c = K.T @ y_star + reduced
    qp = QuadraticProgram()
    qp.constraint_matrix = K
    qp.constraint_lower_bounds = clb
    qp.constraint_upper_bounds = cub
    qp.variable_lower_bounds = lb
    qp.variable_upper_bounds = ub
    if qp_diagonal:
        q = rng.uniform(0.1, 10.0, n)
        c = c - q * x_star  # gradient c + Q x* - K^T y* keeps the planted reduced costs
        qp.set_objective_matrix_diagonal(q)
    qp.objective_vector = c
    qp.problem_name = f"random_sparse_{'qp' if qp_diagonal else 'lp'}_{m}x{n}_seed{seed}"
    qp.planted_primal = x_star
    qp.planted_dual = y_star
    return qp


def multicommodity_flow_lp(num_commodities=64, grid_side=256, num_arcs=520000, pairs_per_commodity=16, seed=0, order="block", commodity_block=8):
    """Config C3: K commodities on a directed grid with random shortcut arcs.

    One conservation row per (commodity, node) with +1 / -1 entries and one joint capacity row per
    arc with K ones: nnz = 3 K A.  Capacities are 1.3 x a feasible Manhattan routing, so the LP is
    feasible; costs are positive so it is bounded.

    The same LP in three variable / constraint orders (a permutation; costs are drawn per (arc,
    commodity) so the instance is identical):
    order="commodity": the textbook block-angular layout - x[k, a] at column k A + a, conservation
      rows at k N + node, then the A linking capacity rows.  A contiguous column block is a set of
      commodities: its conservation rows are private to it and only the capacity rows link blocks.
    order="arc": x at a K + k, conservation rows at node K + k.  A capacity row's K entries and the
      K conservation rows of a node then sit in K consecutive positions (coalesced gathers), but a
      contiguous column block is a set of arcs and touches nearly every row.
    order="block" (default): block-angular with blocks of `commodity_block` commodities, arc-major
      inside a block: block b holds commodities [b B, (b+1) B), x[k, a] at column b A B + a B + (k - b B),
      its conservation rows at b N B + node B + (k - b B), then the capacity rows.  Keeps runs of B
      consecutive entries (B = 8: one 64-byte run) and shards into K / B independent blocks."""
    rng = np.random.Generator(np.random.PCG64(seed))
    side = grid_side
    N = side * side
    node = lambda r, c: r * side + c
    rr, cc = np.meshgrid(np.arange(side), np.arange(side), indexing="ij")
    tails, heads = [], []
    # grid arcs in 4 directions (no wrap)
    for dr, dc in ((0, 1), (0, -1), (1, 0), (-1, 0)):
        ok = (rr + dr >= 0) & (rr + dr < side) & (cc + dc >= 0) & (cc + dc < side)
        tails.append(node(rr[ok], cc[ok]))
        heads.append(node(rr[ok] + dr, cc[ok] + dc))
    tails = np.concatenate(tails)
    heads = np.concatenate(heads)
    grid_arcs = tails.size
    extra = max(0, num_arcs - grid_arcs)
    et = rng.integers(0, N, extra)
    eh = rng.integers(0, N, extra)
    eh = np.where(eh == et, (eh + 1) % N, eh)
    tails = np.concatenate([tails, et])  # all grid arcs are kept: the planted routing uses them
    heads = np.concatenate([heads, eh])
    A = tails.size
    Kc = num_commodities
    # arc id lookup for grid moves: (node, direction) -> arc index
    arc_of = {}
    dirs = ((0, 1), (0, -1), (1, 0), (-1, 0))
    start = 0
    lookup = np.full((4, N), -1, dtype=np.int64)
    for d, (dr, dc) in enumerate(dirs):
        ok = (rr + dr >= 0) & (rr + dr < side) & (cc + dc >= 0) & (cc + dc < side)
        cnt = int(ok.sum())
        lookup[d, node(rr[ok], cc[ok])] = np.arange(start, start + cnt)
        start += cnt
    del arc_of
    # demands and a feasible routing
    demand = np.zeros((N, Kc))
    load = np.zeros(A)
    for k in range(Kc):
        for _ in range(pairs_per_commodity):
            s, t = rng.integers(0, N, 2)
            if s == t:
                continue
            d = rng.uniform(1.0, 10.0)
            demand[s, k] += d
            demand[t, k] -= d
            r, c = divmod(int(s), side)
            tr, tc = divmod(int(t), side)
            while c != tc:
                dirn = 0 if tc > c else 1
                load[lookup[dirn, node(r, c)]] += d
                c += 1 if tc > c else -1
            while r != tr:
                dirn = 2 if tr > r else 3
                load[lookup[dirn, node(r, c)]] += d
                r += 1 if tr > r else -1
    capacity = 1.3 * load + 1.0
    n = A * Kc
    m = N * Kc + A
    # column of (a, k): rows (tail, k) -> +1, (head, k) -> -1, capacity row of a -> +1
    if order == "arc":
        a_idx = np.repeat(np.arange(A, dtype=np.int64), Kc)
        k_idx = np.tile(np.arange(Kc, dtype=np.int64), A)
        r_tail = tails[a_idx] * Kc + k_idx
        r_head = heads[a_idx] * Kc + k_idx
    elif order == "commodity":
        k_idx = np.repeat(np.arange(Kc, dtype=np.int64), A)
        a_idx = np.tile(np.arange(A, dtype=np.int64), Kc)
        r_tail = k_idx * N + tails[a_idx]
        r_head = k_idx * N + heads[a_idx]
    elif order == "block":
        B = commodity_block
        if Kc % B != 0:
            raise ValueError("commodity_block must divide num_commodities")
        b_idx = np.repeat(np.arange(Kc // B, dtype=np.int64), A * B)
        a_idx = np.tile(np.repeat(np.arange(A, dtype=np.int64), B), Kc // B)
        l_idx = np.tile(np.arange(B, dtype=np.int64), A * (Kc // B))
        k_idx = b_idx * B + l_idx
        r_tail = b_idx * (N * B) + tails[a_idx] * B + l_idx
        r_head = b_idx * (N * B) + heads[a_idx] * B + l_idx
    else:
        raise ValueError(order)
    r_cap = N * Kc + a_idx
    rows = np.stack([r_tail, r_head, r_cap], axis=1)
    vals = np.stack([np.ones(n), -np.ones(n), np.ones(n)], axis=1)
    by_row = np.argsort(rows, axis=1, kind="stable")
    rows = np.take_along_axis(rows, by_row, axis=1)
    vals = np.take_along_axis(vals, by_row, axis=1)
    K = _csc_from_columns(m, n, rows, vals, 3)
    qp = QuadraticProgram()
    qp.constraint_matrix = K
    if order == "arc":
        rhs = demand.reshape(-1)  # (node, k)
    elif order == "commodity":
        rhs = demand.T.reshape(-1)  # (k, node)
    else:
        rhs = demand.reshape(N, Kc // commodity_block, commodity_block).transpose(1, 0, 2).reshape(-1)  # (block, node, k in block)
    qp.constraint_lower_bounds = np.concatenate([rhs, np.full(A, -INF)])
    qp.constraint_upper_bounds = np.concatenate([rhs, capacity])
    qp.variable_lower_bounds = np.zeros(n)
    qp.variable_upper_bounds = np.full(n, INF)
    cost = rng.uniform(1.0, 10.0, (A, Kc))  # per (arc, commodity): the same instance in every order
    qp.objective_vector = cost[a_idx, k_idx]
    qp.problem_name = f"multicommodity_flow_K{Kc}_N{N}_A{A}_seed{seed}_{order}"
    return qp


def transportation_lp(num_sources, num_sinks, seed=0, first_source=0, last_source=None, local_rows=False):
    """Config C5: dense bipartite transportation problem; x[s, d] at column s * D + d.

    Rows: supply s (sum_d x <= supply_s), then demand S + d (sum_s x = demand_d); every
    entry of K is 1 and nnz = 2 n.  `first_source` / `last_source` restrict the columns to
    a block of sources (the per-GPU column shard).  local_rows=False keeps all S + D rows;
    local_rows=True keeps only the rows the block touches, in the sharded solver's local order:
    the D demand rows (shared by every block) first, then the block's own supply rows (private)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    S, D = num_sources, num_sinks
    demand = rng.uniform(1.0, 100.0, D)
    supply = rng.uniform(1.0, 100.0, S)
    supply *= 1.05 * demand.sum() / supply.sum()
    last_source = S if last_source is None else last_source
    block = last_source - first_source
    n = block * D
    # Row pairs of every column, written in place (no index temporaries of the matrix's size).
    rows = np.empty((block, D, 2), dtype=np.int32 if S + D < 2 ** 31 else np.int64)
    sources = np.arange(first_source, last_source, dtype=rows.dtype)
    sinks = np.arange(D, dtype=rows.dtype)
    if local_rows:
        rows[:, :, 0] = sinks[None, :]
        rows[:, :, 1] = (D + (sources - first_source))[:, None]
        num_rows = D + block
    else:
        rows[:, :, 0] = sources[:, None]
        rows[:, :, 1] = (S + sinks)[None, :]
        num_rows = S + D
    rows = rows.reshape(n, 2)
    vals = np.ones((n, 2))
    K = _csc_from_columns(num_rows, n, rows, vals, 2)
    # The block's slice of ONE cost stream of S D draws (one 64-bit output per double), so that every
    # sharding of the same (S, D, seed) is the same LP and no rank generates the other blocks' costs.
    cost = np.random.Generator(np.random.PCG64([seed, 1]).advance(first_source * D)).uniform(1.0, 100.0, n)
    qp = QuadraticProgram()
    qp.constraint_matrix = K
    if local_rows:
        qp.constraint_lower_bounds = np.concatenate([demand, np.full(block, -INF)])
        qp.constraint_upper_bounds = np.concatenate([demand, supply[first_source:last_source]])
    else:
        qp.constraint_lower_bounds = np.concatenate([np.full(S, -INF), demand])
        qp.constraint_upper_bounds = np.concatenate([supply, demand])
    qp.variable_lower_bounds = np.zeros(n)
    qp.variable_upper_bounds = np.full(n, INF)
    qp.objective_vector = cost
    qp.problem_name = f"transportation_{S}x{D}_seed{seed}"
    return qp

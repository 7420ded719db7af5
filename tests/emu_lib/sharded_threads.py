"""TEST INFRASTRUCTURE ONLY.  Run by tests/test_full_emulation.py in a child process with PDLP_B200_LIBRARY = the emulated
library and PDLP_B200_NCCL_LIBRARY = the fake NCCL: a column-sharded solve whose ranks are host threads of this process
(the emulated peer memory is plain memory, tests/emu_lib/include/cuda_runtime.h), checked against the one-rank solve of
the same library.      python sharded_threads.py <world> <peer|nccl> [random|flow]      prints one JSON line.

random: a random sparse LP, nearly every row shared by all ranks.  flow: a small multi-commodity flow LP in the benchmark's
block-angular order (config 3 in miniature): conservation rows private to a rank, capacity rows shared - the layout that makes
the fused peer pass split its grid between pushing and private CTAs and signal from inside the kernel."""
import json
import os
import sys
import threading

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import fixtures  # noqa: E402
import pdlp_b200  # noqa: E402
from pdlp_b200 import sharding  # noqa: E402
import scipy.sparse as sp  # noqa: E402


def solve_sharded(qp, params, world, corrupt_rank=-1):
    K = sp.csc_matrix(qp.constraint_matrix)
    starts = sharding.column_partition(K.indptr, world)
    sharing = sharding.RowSharing.from_matrix(K, starts)
    uid = sharding.unique_id()
    results, errors = [None] * world, []

    def rank_main(rank):
        try:
            lo, hi = int(starts[rank]), int(starts[rank + 1])
            block = sharding.take_columns(qp, lo, hi, sharing.local_rows(rank))
            if rank == corrupt_rank:  # a row index outside the block: BuildDeviceMatrix rejects it on this rank only
                bad = sp.csc_matrix(block.constraint_matrix).copy()
                bad.indices[0] = bad.shape[0] + 7
                block.constraint_matrix = bad
            shard = sharding.ColumnShard(rank, world, int(starts[-1]), lo, uid, int(sharing.shared.size))
            r = pdlp_b200.primal_dual_hybrid_gradient(block, params, device=rank, shard=shard)
            results[rank] = (r, pdlp_b200.last_solve_counters())
        except Exception as e:  # noqa: BLE001
            errors.append(f"rank {rank}: {e!r}")

    threads = [threading.Thread(target=rank_main, args=(r,)) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    if corrupt_rank >= 0:
        return errors, [r is not None for r in results]
    if errors:
        raise RuntimeError("; ".join(errors))
    x = np.concatenate([r[0].primal_solution for r in results])
    y = sharing.gather([r[0].dual_solution for r in results])
    return results, x, y, sharing


def main():
    world, mode = int(sys.argv[1]), sys.argv[2]
    if mode == "nccl":
        os.environ["PDLP_B200_PEER_EXCHANGE"] = "0"
    problem = sys.argv[3] if len(sys.argv) > 3 else "random"
    if problem == "bad_block":
        # One rank's block is not a valid CSC matrix.  Every rank must come back with an error - the others must not be
        # left waiting in a collective for a rank that has given up (PreprocessSolver agrees on a failed build).
        qp = fixtures.random_lp(40, 100, 4, 21)
        errors, returned = solve_sharded(qp, pdlp_b200.PrimalDualHybridGradientParams(), world, corrupt_rank=world - 1)
        print(json.dumps({"world": world, "problem": problem, "errors": errors, "returned_a_result": returned}))
        return
    if problem == "flow":
        from pdlp_b200 import synthetic

        qp = synthetic.multicommodity_flow_lp(num_commodities=8, grid_side=6, num_arcs=120, pairs_per_commodity=3, commodity_block=1, seed=3)  # 408 x 960
        eps = 1e-4  # 448 iterations on the oracle
    elif problem == "qp_mp":  # a diagonal QP under the Malitsky-Pock rule with feasibility-polishing-free defaults
        qp = fixtures.random_lp(40, 100, 4, 21, qp_diagonal=True)
        eps = 1e-6
    else:
        qp = fixtures.random_lp(40, 100, 4, 21)  # 448 iterations on the oracle
        eps = 1e-6
    params = pdlp_b200.PrimalDualHybridGradientParams()
    if problem == "qp_mp":
        params.linesearch_rule = params.MALITSKY_POCK_LINESEARCH_RULE
    params.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = eps
    params.termination_criteria.simple_optimality_criteria.eps_optimal_relative = eps
    params.termination_criteria.iteration_limit = 20000
    single = pdlp_b200.primal_dual_hybrid_gradient(qp, params, device=0)
    results, x, y, sharing = solve_sharded(qp, params, world)
    logs = [r[0].solve_log for r in results]
    c = np.asarray(qp.objective_vector)
    q = None if qp.objective_matrix is None else np.asarray(qp.objective_matrix)
    objective = lambda v: float(c @ v + (0.0 if q is None else 0.5 * (q * v) @ v))  # noqa: E731
    out = {
        "world": world, "mode": mode, "problem": problem, "eps": eps,
        "private_rows": [int(p.size) for p in sharing.private],
        "single_reason": int(single.solve_log.termination_reason), "single_iterations": int(single.solve_log.iteration_count),
        "reasons": [int(l.termination_reason) for l in logs], "iterations": [int(l.iteration_count) for l in logs],
        "objective_single": objective(single.primal_solution), "objective_sharded": objective(x),
        "x_distance": float(np.linalg.norm(x - single.primal_solution) / max(1.0, np.linalg.norm(single.primal_solution))),
        "y_distance": float(np.linalg.norm(y - single.dual_solution) / max(1.0, np.linalg.norm(single.dual_solution))),
        "peer_exchanges": [int(r[1]["peer_exchanges"]) for r in results], "collectives": [int(r[1]["collectives"]) for r in results],
        "shared_rows": int(sharing.shared.size), "rows": int(qp.constraint_matrix.shape[0]),
        "duals_identical_on_shared_rows": bool(all(np.array_equal(results[0][0].dual_solution[:sharing.shared.size], r[0].dual_solution[:sharing.shared.size]) for r in results)),
    }
    print(json.dumps(out))


if __name__ == "__main__":
    main()

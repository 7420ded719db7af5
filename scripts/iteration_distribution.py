#!/usr/bin/env python
"""Iteration counts of one backend over seeds x 1-ulp perturbations of the data (evidence for the
iteration-count parity criterion; tests/test_gpu_iteration_parity.py asserts on the same numbers).

    python scripts/iteration_distribution.py oracle|device OUT.json [--seeds 20] [--perturb 6] [--m 400 --n 800] [--eps 1e-6]

Perturbation k > 0 multiplies every objective coefficient by (1 + u_k), u_k ~ 1e-15 N(0, 1) (seeded
by k): a relative change of a few ulp of the DATA.  k = 0 is the unperturbed problem.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import conftest  # noqa: E402,F401
import fixtures  # noqa: E402
import pdlp_b200  # noqa: E402


def perturbed(qp, k):
    if k == 0:
        return qp
    q = qp.__class__.__new__(qp.__class__)
    q.__dict__.update(qp.__dict__)
    rng = np.random.default_rng(1000 + k)
    q.objective_vector = qp.objective_vector * (1 + 1e-15 * rng.standard_normal(len(qp.objective_vector)))
    return q


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("backend", choices=["oracle", "device"])
    ap.add_argument("out")
    ap.add_argument("--seeds", type=int, default=20)
    ap.add_argument("--perturb", type=int, default=6)
    ap.add_argument("--m", type=int, default=400)
    ap.add_argument("--n", type=int, default=800)
    ap.add_argument("--eps", type=float, default=1e-6)
    ap.add_argument("--qp", action="store_true")
    args = ap.parse_args()
    p = pdlp_b200.PrimalDualHybridGradientParams()
    p.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = args.eps
    p.termination_criteria.simple_optimality_criteria.eps_optimal_relative = args.eps
    if args.backend == "oracle":
        import oracle_binding as ob
        solve = lambda qp: ob.solve(qp, p).solve_log_dict  # noqa: E731
        get = lambda log: (log["iteration_count"], log["termination_reason"])  # noqa: E731
    else:
        solve = lambda qp: pdlp_b200.primal_dual_hybrid_gradient(qp, p, device=0).solve_log  # noqa: E731
        get = lambda log: (int(log.iteration_count), int(log.termination_reason))  # noqa: E731
    out = {"backend": args.backend, "m": args.m, "n": args.n, "eps": args.eps, "qp": args.qp, "counts": {}}
    t0 = time.time()
    for seed in range(args.seeds):
        qp = fixtures.random_lp(args.m, args.n, 5, seed, qp_diagonal=args.qp)
        row = []
        for k in range(args.perturb + 1):
            its, reason = get(solve(perturbed(qp, k)))
            row.append(its if reason == 1 else -its)
        out["counts"][str(seed)] = row
        print(seed, row, f"{time.time() - t0:.0f}s", flush=True)
    with open(args.out, "w") as f:
        json.dump(out, f)


if __name__ == "__main__":
    main()

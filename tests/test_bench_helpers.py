"""bench.py's own checkers against the CPU oracle (no GPU): the numpy restatement of
ComputeConvergenceInformation on the transportation structure, which stands in for the oracle on the
c5 lines whose matrix never exists on one host."""
import os
import sys
import threading

import numpy as np

import oracle_binding as ob

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def _point(n, m, seed):
    rng = np.random.default_rng(seed)
    x = np.maximum(rng.normal(0.5, 1.0, n), 0.0)  # projected onto x >= 0 like every PDHG iterate
    y = rng.normal(0.0, 30.0, m)
    return x, y


def _feasible_signs(qp, y):
    """Duals of rows without a lower bound projected onto y <= 0, like every PDHG iterate."""
    return np.where(np.isfinite(qp.constraint_lower_bounds), y, np.minimum(y, 0.0))


def _oracle(qp, x, y):
    prob = ob.Problem(qp, num_threads=1, num_shards=4)
    st = prob.stats()
    ci = prob.convergence_info(bench.oracle_params_text("c5", 1), np.ones(len(x)), np.ones(len(y)), x, y)
    return ci, bench.relative_kkt(ci, 1e-4, st["combined_bounds_l2_norm"], st["objective_vector_l2_norm"])


def _close(a, b):
    return abs(a - b) <= 1e-11 * (1.0 + abs(b))


def test_transportation_checker_matches_the_oracle_on_the_whole_problem():
    synthetic = bench.synthetic_module(False)
    qp = synthetic.transportation_lp(13, 9, seed=3)
    x, y = _point(13 * 9, 13 + 9, 5)
    y = _feasible_signs(qp, y)
    ci, want = _oracle(qp, x, y)
    got = bench.transportation_parity(qp, x, y, False, ci["primal_objective"], 7, lambda v: v)
    assert _close(got["checker_primal_objective"], ci["primal_objective"]) and _close(got["checker_dual_objective"], ci["dual_objective"])
    for k, v in want.items():
        assert _close(got["relative_kkt_over_eps"][k], v), (k, got["relative_kkt_over_eps"][k], v)
    assert got["ok"] is False  # a random point is not optimal
    assert got["objective_vs_checker_rel"] <= 1e-12


def test_transportation_checker_over_two_blocks_adds_up_to_the_whole():
    synthetic = bench.synthetic_module(False)
    S, D, cut = 12, 7, 5
    whole = synthetic.transportation_lp(S, D, seed=2)
    x, y = _point(S * D, S + D, 11)  # whole layout: supply rows, then sink rows
    y = _feasible_signs(whole, y)
    ci, want = _oracle(whole, x, y)
    blocks = [synthetic.transportation_lp(S, D, 2, 0, cut, local_rows=True), synthetic.transportation_lp(S, D, 2, cut, S, local_rows=True)]
    assert np.array_equal(np.concatenate([b.objective_vector for b in blocks]), whole.objective_vector)  # one cost stream for every sharding
    xs = [x[:cut * D], x[cut * D:]]
    ys = [np.concatenate([y[S:], y[:cut]]), np.concatenate([y[S:], y[cut:S]])]  # local layout: sink rows, then the block's supply rows
    barrier = threading.Barrier(2)
    lock = threading.Lock()
    pending = {}

    def reduce_sum(rank):
        def f(v):
            with lock:
                pending.setdefault("in", []).append(np.array(v, dtype=float))
            barrier.wait()
            total = sum(pending["in"])
            barrier.wait()
            if rank == 0:
                pending["in"] = []
            barrier.wait()
            return total
        return f

    out = [None, None]

    def run(rank):
        out[rank] = bench.transportation_parity(blocks[rank], xs[rank], ys[rank], True, ci["primal_objective"], 7, reduce_sum(rank))

    threads = [threading.Thread(target=run, args=(r,)) for r in range(2)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for got in out:
        assert _close(got["checker_primal_objective"], ci["primal_objective"]) and _close(got["checker_dual_objective"], ci["dual_objective"])
        for k, v in want.items():
            assert _close(got["relative_kkt_over_eps"][k], v), (k, got["relative_kkt_over_eps"][k], v)


def test_n1_reference_is_keyed_by_workload_and_scale(tmp_path, monkeypatch):
    """An N = 1 solve leaves (iterations, objective) behind; an N > 1 solve of the SAME workload and scale compares with
    it (two eps-optimal points of one LP may differ by 4 eps), another scale finds nothing to compare with."""
    import tempfile
    monkeypatch.setattr(tempfile, "tempdir", str(tmp_path))
    out = {}
    assert bench.compare_with_n1(out, "c5", 8.0, 1, 9024, 1143586.5, 1e-4) and "vs_n1" not in out
    out = {}
    assert bench.compare_with_n1(out, "c5", 8.0, 8, 9600, 1143530.3, 1e-4)
    assert out["vs_n1"]["iterations_n1"] == 9024 and abs(out["vs_n1"]["iteration_ratio"] - 9600 / 9024) < 1e-12
    assert 4e-5 < out["vs_n1"]["objective_rel_diff"] < 6e-5
    out = {}
    assert not bench.compare_with_n1(out, "c5", 8.0, 8, 9600, 1143586.5 * (1 + 1e-3), 1e-4)  # not the same optimum
    out = {}
    assert bench.compare_with_n1(out, "c5", 2.0, 8, 100, 1.0, 1e-4) and "vs_n1" not in out  # no N = 1 run of that scale on this box

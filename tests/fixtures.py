"""The reference's fixture problems, re-typed from
/root/reference/ortools/pdlp/test_util.cc:35-259 (expected optima in test_util.h:33-202).
Each builder returns a fresh pdlp_b200.QuadraticProgram."""
import numpy as np
import scipy.sparse as sp

import pdlp_b200

INF = np.inf


def _qp(n, m, triplets, c, clb, cub, vlb, vub, offset=0.0, q=None):
    qp = pdlp_b200.QuadraticProgram()
    qp.resize_and_initialize(n, m)
    if triplets:
        r, cidx, v = zip(*triplets)
        qp.constraint_matrix = sp.csc_matrix((np.array(v, dtype=float), (np.array(r), np.array(cidx))), shape=(m, n))
    qp.objective_vector = np.array(c, dtype=float)
    qp.constraint_lower_bounds = np.array(clb, dtype=float)
    qp.constraint_upper_bounds = np.array(cub, dtype=float)
    qp.variable_lower_bounds = np.array(vlb, dtype=float)
    qp.variable_upper_bounds = np.array(vub, dtype=float)
    qp.objective_offset = offset
    if q is not None:
        qp.set_objective_matrix_diagonal(np.array(q, dtype=float))
    return qp


def make_test_lp():
    # optimum -34: x = [-1, 8, 1, 2.5], y = [-2, 0, 2.375, 2/3], rc = [0, 0, 0, 5 2/3]... (test_util.h:45-48)
    return _qp(4, 4, [(0, 0, 2), (0, 1, 1), (0, 2, 1), (0, 3, 2), (1, 0, 1), (1, 2, 1), (2, 0, 4), (3, 2, 1.5), (3, 3, -1)],
               [5.5, -2, -1, 1], [12, -INF, -4, -1], [12, 7, INF, 1], [-INF, -2, -INF, 2.5], [INF, INF, 6, 3.5], offset=-14)


def tiny_lp():
    # optimum -1: x = [1, 0, 6, 2], y = [0.5, 4, 0], rc = [0, 1.5, -3.5, 0] (test_util.h:63-66)
    return _qp(4, 3, [(0, 0, 2), (0, 1, 1), (0, 2, 1), (0, 3, 2), (1, 0, 1), (1, 2, 1), (2, 2, 1), (2, 3, -1)],
               [5, 2, 1, 1], [12, 7, 1], [12, INF, INF], [0, 0, 0, 0], [2, 4, 6, 3], offset=-14)


def correlation_clustering_lp():
    # optimum 1: x = [1, 1, 0, 1, 0, 0], multiple optimal duals (test_util.h:70-88)
    return _qp(6, 3, [(0, 1, -1), (0, 2, 1), (0, 5, -1), (1, 3, -1), (1, 4, 1), (1, 5, -1), (2, 0, -1), (2, 1, -1), (2, 3, 1)],
               [-1, -1, 1, -1, 1, -1], [-1, -1, -1], [INF] * 3, [0] * 6, [1] * 6, offset=4)


def correlation_clustering_star_lp():
    # optimum 1.5: x = [0.5, 0.5, 0.5, 0, 0, 0], y = [0.5, 0.5, 0.5] (test_util.h:99-106)
    return _qp(6, 3, [(0, 0, -1), (0, 1, -1), (0, 3, 1), (1, 0, -1), (1, 2, -1), (1, 4, 1), (2, 1, -1), (2, 2, -1), (2, 5, 1)],
               [-1, -1, -1, 1, 1, 1], [-1, -1, -1], [INF] * 3, [0] * 6, [1] * 6, offset=3)


def diagonal_qp1():
    # optimum: x = [1, 0], y = [-1], objective 6, rc = [4, 0] (test_util.h:118-121)
    return _qp(2, 1, [(0, 0, 1), (0, 1, 1)], [-1, -1], [-INF], [1], [1, -2], [2, 4], offset=5, q=[4.0, 1.0])


def diagonal_qp2():
    # optimum: x = [3, 1], y = [0], objective -5 (test_util.h:131-133)
    return _qp(2, 1, [(0, 0, 1), (0, 1, -1)], [-3, -1], [2], [2], [0, 0], [INF, INF], offset=0, q=[1.0, 1.0])


def diagonal_qp3():
    # optimum: x = [2, 0, 1], y = [-1, 1], objective 2 (test_util.h:145-147)
    return _qp(3, 2, [(0, 0, 1), (0, 2, -1), (1, 0, 2)], [1, 0, -1], [1, 4], [1, 4], [0, 0, 0], [INF, INF, INF], offset=0, q=[0.0, 1.0, 2.0])


def small_invalid_problem_lp():
    return _qp(2, 1, [(0, 0, 1), (0, 1, -1)], [1, 1], [2.0], [1.0], [0, 0], [INF, INF])


def small_inconsistent_variable_bounds_lp():
    return _qp(2, 1, [(0, 0, 1), (0, 1, -1)], [1, 1], [-INF], [1.0], [2.0, 0.0], [1.0, INF])


def small_primal_infeasible_lp():
    return _qp(2, 2, [(0, 0, 1), (0, 1, -1), (1, 0, -1), (1, 1, 1)], [1, 1], [-INF, -INF], [1.0, -2.0], [0, 0], [INF, INF])


def small_dual_infeasible_lp():
    qp = small_primal_infeasible_lp()
    qp.constraint_upper_bounds[1] = 2.0
    qp.objective_vector = -qp.objective_vector
    return qp


def small_primal_dual_infeasible_lp():
    qp = small_primal_infeasible_lp()
    qp.objective_vector = -qp.objective_vector
    return qp


def small_initialization_lp():
    return _qp(2, 2, [(0, 0, 1), (0, 1, 1), (1, 0, 1), (1, 1, 2)], [-4.0, 0.0], [-INF, -INF], [2.0, 2.0], [0.5, 0.5], [2.0, 2.0])


def lp_without_constraints():
    return _qp(2, 0, [], [4.0, 0.0], [], [], [0.0, -INF], [INF, 0.0])


def random_lp(m, n, nnz_per_col, seed, equality_fraction=0.5, qp_diagonal=False):
    """C2-style synthetic LP: feasible and bounded by construction."""
    from pdlp_b200 import synthetic
    return synthetic.random_sparse_lp(m, n, nnz_per_col, seed, equality_fraction, qp_diagonal)

"""ortools/pdlp/termination_test.cc re-typed (the IterationStats literals are the reference's own
text protos).  The predicates are host code on both sides: the oracle's and the product's
(pdlp_b200_check_termination in the C-ABI library) are both checked on the CPU box."""
import math

import numpy as np
import pytest
from google.protobuf import text_format

import oracle_binding as ob
import pdlp_b200

P = pdlp_b200.PrimalDualHybridGradientParams
TR = pdlp_b200.TerminationReason
PT = pdlp_b200.PointType
INF = math.inf
TEST_LP_NORMS = [math.sqrt(36.25), math.sqrt(210.0), 5.5, 12.0]  # :63-68 (l2 objective, l2 bounds, linf objective, linf bounds)
ZERO_NORMS = [0.0, 0.0, 0.0, 0.0]
NORMS = [1, 2, 3]  # OPTIMALITY_NORM_L_INF, L2, L_INF_COMPONENTWISE


@pytest.fixture(params=["oracle", "product"])
def check(request):
    return ob.check_termination if request.param == "oracle" else pdlp_b200.check_termination


def stats(text=""):
    return text_format.Parse(text, pdlp_b200.IterationStats())


def criteria(text, norm=None):
    p = P()
    text_format.Parse(text, p.termination_criteria)
    if norm is not None:
        p.termination_criteria.optimality_norm = norm
    return p


def simple():  # SimpleTerminationTest (:78-88)
    return criteria("time_sec_limit: 1.0 kkt_matrix_pass_limit: 2000 iteration_limit: 10")


def iterate(norm):  # IterateTerminationTest (:90-107)
    return criteria("""simple_optimality_criteria { eps_optimal_absolute: 1.0e-4 eps_optimal_relative: 1.0e-4 }
                       eps_primal_infeasible: 1.0e-6 eps_dual_infeasible: 1.0e-6 time_sec_limit: 1.0 kkt_matrix_pass_limit: 2000 iteration_limit: 10""", norm)


def detailed_relative(norm):  # :109-126
    return criteria("""detailed_optimality_criteria { eps_optimal_primal_residual_absolute: 0.0 eps_optimal_primal_residual_relative: 1.0e-4
                       eps_optimal_dual_residual_absolute: 0.0 eps_optimal_dual_residual_relative: 1.0e-4
                       eps_optimal_objective_gap_absolute: 0.0 eps_optimal_objective_gap_relative: 1.0e-4 }""", norm)


def detailed_absolute(norm):  # :128-145
    return criteria("""detailed_optimality_criteria { eps_optimal_primal_residual_absolute: 1.0e-4 eps_optimal_primal_residual_relative: 0.0
                       eps_optimal_dual_residual_absolute: 1.0e-4 eps_optimal_dual_residual_relative: 0.0
                       eps_optimal_objective_gap_absolute: 1.0e-4 eps_optimal_objective_gap_relative: 0.0 }""", norm)


def conv(po, do, linf_p, linf_d, l2_p, l2_d, cw_p, cw_d, typed=True):
    return stats(f"""convergence_information {{ primal_objective: {po} dual_objective: {do} l_inf_primal_residual: {linf_p}
        l_inf_dual_residual: {linf_d} l2_primal_residual: {l2_p} l2_dual_residual: {l2_d} l_inf_componentwise_primal_residual: {cw_p}
        l_inf_componentwise_dual_residual: {cw_d} {'candidate_type: POINT_TYPE_CURRENT_ITERATE' if typed else ''} }}""")


OPTIMAL_CURRENT = (TR.TERMINATION_REASON_OPTIMAL, PT.POINT_TYPE_CURRENT_ITERATE)
NOTHING = (0, 0)


def result(check, p, s, norms=TEST_LP_NORMS, **kw):
    r = check(p, s, norms, **kw)
    return (r[0], r[1])


# ---- detailed criteria (:215-403) ----
REL_NEAR = (1.00019, 1.0, 11.0e-4, 5.4e-4, 14.0e-4, 6.0e-4, 9.0e-5, 9.0e-5)
ABS_NEAR = (1.00009, 1.0, 9.0e-5, 9.0e-5, 9.0e-5, 9.0e-5, 0.0, 0.0)


@pytest.mark.parametrize("norm", NORMS)
def test_detailed_relative(check, norm):
    p = detailed_relative(norm)
    assert result(check, p, conv(*REL_NEAR)) == OPTIMAL_CURRENT
    assert result(check, p, conv(1.00021, *REL_NEAR[1:])) == NOTHING  # excessive gap
    assert result(check, p, conv(1.00019, 1.0, 13.0e-4, 5.4e-4, 15.0e-4, 6.0e-4, 1.1e-4, 9.0e-5)) == NOTHING  # primal residual
    assert result(check, p, conv(1.00019, 1.0, 11.0e-4, 5.6e-4, 14.0e-4, 7.0e-4, 9.0e-5, 1.1e-4)) == NOTHING  # dual residual


@pytest.mark.parametrize("norm", NORMS)
def test_detailed_absolute(check, norm):
    p = detailed_absolute(norm)
    assert result(check, p, conv(*ABS_NEAR)) == OPTIMAL_CURRENT
    assert result(check, p, conv(1.00011, *ABS_NEAR[1:])) == NOTHING
    if norm != 3:  # the componentwise norm ignores the absolute tolerance (termination.cc:79-87)
        assert result(check, p, conv(1.00009, 1.0, 1.1e-4, 9.0e-5, 1.1e-4, 9.0e-5, 0.0, 0.0)) == NOTHING
        assert result(check, p, conv(1.00009, 1.0, 9.0e-5, 1.1e-4, 9.0e-5, 1.1e-4, 0.0, 0.0)) == NOTHING


# ---- simple criteria (:417-477) ----
def test_simple_criteria(check):
    p = simple()
    assert result(check, p, stats(), which="simple") == NOTHING
    flag = np.ones(1, dtype=np.int32)
    assert result(check, p, stats(), which="simple", interrupt=flag) == (TR.TERMINATION_REASON_INTERRUPTED_BY_USER, PT.POINT_TYPE_NONE)
    assert result(check, p, stats(), which="simple", interrupt=np.zeros(1, dtype=np.int32)) == NOTHING
    assert result(check, p, stats("cumulative_time_sec: 100.0"), which="simple") == (TR.TERMINATION_REASON_TIME_LIMIT, PT.POINT_TYPE_NONE)
    assert result(check, p, stats("cumulative_kkt_matrix_passes: 2500"), which="simple") == (TR.TERMINATION_REASON_KKT_MATRIX_PASS_LIMIT, PT.POINT_TYPE_NONE)
    assert result(check, p, stats("iteration_number: 20"), which="simple") == (TR.TERMINATION_REASON_ITERATION_LIMIT, PT.POINT_TYPE_NONE)


# ---- iterate criteria (:405-825) ----
@pytest.mark.parametrize("norm", NORMS)
def test_iterate_criteria(check, norm):
    p = iterate(norm)
    assert result(check, p, stats("convergence_information { primal_objective: 50.0 dual_objective: -50.0 }")) == NOTHING
    assert result(check, p, stats()) == NOTHING
    assert result(check, p, stats(), force_numerical_termination=True) == (TR.TERMINATION_REASON_NUMERICAL_ERROR, PT.POINT_TYPE_NONE)
    # rays
    s = stats("infeasibility_information: { dual_ray_objective: 1.0 max_dual_ray_infeasibility: 1.0e-16 candidate_type: POINT_TYPE_ITERATE_DIFFERENCE }")
    assert result(check, p, s) == (TR.TERMINATION_REASON_PRIMAL_INFEASIBLE, PT.POINT_TYPE_ITERATE_DIFFERENCE)
    assert result(check, p, stats("infeasibility_information: { dual_ray_objective: 1.0 max_dual_ray_infeasibility: 1.0e-5 }")) == NOTHING
    assert result(check, p, stats("infeasibility_information: { dual_ray_objective: -1.0 max_dual_ray_infeasibility: 0.0 }")) == NOTHING
    assert result(check, p, stats("infeasibility_information: { dual_ray_objective: 0.0 max_dual_ray_infeasibility: 0.0 }")) == NOTHING
    s = stats("infeasibility_information: { primal_ray_linear_objective: -1.0 max_primal_ray_infeasibility: 1.0e-16 candidate_type: POINT_TYPE_AVERAGE_ITERATE }")
    assert result(check, p, s) == (TR.TERMINATION_REASON_DUAL_INFEASIBLE, PT.POINT_TYPE_AVERAGE_ITERATE)
    assert result(check, p, stats("infeasibility_information: { primal_ray_linear_objective: -1.0 max_primal_ray_infeasibility: 1.0e-5 }")) == NOTHING
    assert result(check, p, stats("infeasibility_information: { primal_ray_linear_objective: 1.0 max_primal_ray_infeasibility: 0.0 }")) == NOTHING
    assert result(check, p, stats("infeasibility_information: { primal_ray_linear_objective: 0.0 max_primal_ray_infeasibility: 0.0 }")) == NOTHING
    # optimality
    zero = (1.0, 1.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0)
    assert result(check, p, conv(*zero)) == OPTIMAL_CURRENT
    assert result(check, p, conv(*REL_NEAR)) == OPTIMAL_CURRENT
    assert result(check, p, conv(*zero), force_numerical_termination=True) == OPTIMAL_CURRENT  # optimal even with a numerical error
    assert result(check, p, conv(10.0, 1.0, 0, 0, 0, 0, 0, 0, typed=False)) == NOTHING
    assert result(check, p, conv(0, "-inf", 0, 0, 0, 0, 0, 0, typed=False)) == NOTHING
    assert result(check, p, conv(1.0, 1.0, 1.0, 0.0, 1.0, 0.0, 1.0, 0.0, typed=False)) == NOTHING
    assert result(check, p, conv(1.0, 1.0, 0.0, 1.0, 0.0, 1.0, 0.0, 1.0, typed=False)) == NOTHING
    ones = (1.0, 1.0, 1.0, 1.0, 1.0, 1.0, 1.0, 1.0)
    for eps_abs, eps_rel, point in ((INF, 0.0, ones), (0.0, INF, (0.0, 0.0) + ones[2:]), (INF, INF, ones)):
        q = iterate(norm)
        q.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = eps_abs
        q.termination_criteria.simple_optimality_criteria.eps_optimal_relative = eps_rel
        assert result(check, q, conv(*point), norms=ZERO_NORMS) == OPTIMAL_CURRENT
    q = iterate(norm)
    q.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = 0.0
    q.termination_criteria.simple_optimality_criteria.eps_optimal_relative = 0.0
    assert result(check, q, conv(*zero)) == OPTIMAL_CURRENT  # ZeroToleranceZeroError


def test_optimality_norms_differ(check):  # :827-904
    # (L_inf <= 12, L2 <= 14.49, componentwise <= 1), in NORMS order
    expected = {0.5: (True, True, True), 10.0: (True, True, False), 13.0: (False, True, False), 15.0: (False, False, False)}
    for residual, wants in expected.items():
        s = stats(f"""convergence_information {{ primal_objective: 1.0 dual_objective: 1.0 l_inf_primal_residual: {residual}
            l2_primal_residual: {residual} l_inf_componentwise_primal_residual: {residual} candidate_type: POINT_TYPE_CURRENT_ITERATE }}""")
        for norm, want in zip(NORMS, wants):
            p = criteria("simple_optimality_criteria { eps_optimal_relative: 1.0 }", norm)
            assert result(check, p, s) == (OPTIMAL_CURRENT if want else NOTHING), (residual, norm)


# ---- ComputeRelativeResiduals (:933-1160) ----
def relative(check, p):
    s = stats("""convergence_information { primal_objective: 10.0 dual_objective: 5.0 l_inf_primal_residual: 1.0 l2_primal_residual: 1.0
                 l_inf_dual_residual: 1.0 l2_dual_residual: 1.0 }""")
    return list(check(p, s, TEST_LP_NORMS)[2])


def simple_eps(a, r):
    p = P()
    p.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = a
    p.termination_criteria.simple_optimality_criteria.eps_optimal_relative = r
    return p


def with_ratio(rp, rd, rg):
    return [1.0 / (rp + 12.0), 1.0 / (rp + math.sqrt(210.0)), 1.0 / (rd + 5.5), 1.0 / (rd + math.sqrt(36.25)), 5.0 / (rg + 15.0)]


def test_relative_residuals(check):
    assert relative(check, simple_eps(0.0, 1.0e-6)) == with_ratio(0.0, 0.0, 0.0)
    assert relative(check, simple_eps(1.0e-6, 0.0)) == [0.0] * 5
    assert relative(check, simple_eps(1.0e-6, 1.0e-6)) == with_ratio(1.0, 1.0, 1.0)
    assert relative(check, simple_eps(0.0, 0.0)) == with_ratio(1.0, 1.0, 1.0)
    assert relative(check, simple_eps(INF, 1.0e-6)) == [0.0] * 5
    assert relative(check, simple_eps(1.0e-6, INF)) == with_ratio(0.0, 0.0, 0.0)
    assert relative(check, simple_eps(INF, INF)) == with_ratio(1.0, 1.0, 1.0)
    p = criteria("""detailed_optimality_criteria { eps_optimal_primal_residual_absolute: 2.0e-6 eps_optimal_primal_residual_relative: 2.0e-4
                    eps_optimal_dual_residual_absolute: 1.0e-3 eps_optimal_dual_residual_relative: 1.0e-4
                    eps_optimal_objective_gap_absolute: 3.0e-8 eps_optimal_objective_gap_relative: 3.0e-7 }""")
    got = relative(check, p)
    want = with_ratio(0.01, 10.0, 0.1)
    assert all(abs(g - w) <= 4 * np.spacing(w) for g, w in zip(got, want)), (got, want)

"""ortools/pdlp/solvers_proto_validation_test.cc re-typed.  Parameter validation is host code
on both sides, so the oracle's validator and the product's (pdlp_b200_validate_params in the
C-ABI library, no GPU needed) are both checked here on the CPU box."""
import math

import pytest

import oracle_binding as ob
import pdlp_b200

P = pdlp_b200.PrimalDualHybridGradientParams
NAN = math.nan


@pytest.fixture(params=["oracle", "product"])
def validate(request):
    return ob.validate_params if request.param == "oracle" else pdlp_b200.validate_params


def setter(path, value):
    def apply(p):
        obj = p
        parts = path.split(".")
        for part in parts[:-1]:
            obj = getattr(obj, part)
            obj.SetInParent()
        setattr(obj, parts[-1], value)
    return apply


T = "termination_criteria."
BAD = [
    # (field path, bad value)  -> message must name the field (solvers_proto_validation_test.cc:40-678)
    (T + "optimality_norm", 0),
    (T + "eps_optimal_absolute", -1.0), (T + "eps_optimal_absolute", NAN),
    (T + "eps_optimal_relative", -1.0), (T + "eps_optimal_relative", NAN),
    (T + "simple_optimality_criteria.eps_optimal_absolute", -1.0), (T + "simple_optimality_criteria.eps_optimal_absolute", NAN),
    (T + "simple_optimality_criteria.eps_optimal_relative", -1.0), (T + "simple_optimality_criteria.eps_optimal_relative", NAN),
    (T + "detailed_optimality_criteria.eps_optimal_primal_residual_absolute", -1.0), (T + "detailed_optimality_criteria.eps_optimal_primal_residual_absolute", NAN),
    (T + "detailed_optimality_criteria.eps_optimal_primal_residual_relative", -1.0), (T + "detailed_optimality_criteria.eps_optimal_primal_residual_relative", NAN),
    (T + "detailed_optimality_criteria.eps_optimal_dual_residual_absolute", -1.0), (T + "detailed_optimality_criteria.eps_optimal_dual_residual_absolute", NAN),
    (T + "detailed_optimality_criteria.eps_optimal_dual_residual_relative", -1.0), (T + "detailed_optimality_criteria.eps_optimal_dual_residual_relative", NAN),
    (T + "detailed_optimality_criteria.eps_optimal_objective_gap_absolute", -1.0), (T + "detailed_optimality_criteria.eps_optimal_objective_gap_absolute", NAN),
    (T + "detailed_optimality_criteria.eps_optimal_objective_gap_relative", -1.0), (T + "detailed_optimality_criteria.eps_optimal_objective_gap_relative", NAN),
    (T + "eps_primal_infeasible", -1.0), (T + "eps_primal_infeasible", NAN),
    (T + "eps_dual_infeasible", -1.0), (T + "eps_dual_infeasible", NAN),
    (T + "time_sec_limit", -1.0), (T + "time_sec_limit", NAN),
    (T + "iteration_limit", -1),
    (T + "kkt_matrix_pass_limit", -1.0), (T + "kkt_matrix_pass_limit", NAN),
    ("adaptive_linesearch_parameters.step_size_reduction_exponent", 0.0), ("adaptive_linesearch_parameters.step_size_reduction_exponent", 2.0),
    ("adaptive_linesearch_parameters.step_size_reduction_exponent", NAN), ("adaptive_linesearch_parameters.step_size_reduction_exponent", -1.0),
    ("adaptive_linesearch_parameters.step_size_growth_exponent", 0.0), ("adaptive_linesearch_parameters.step_size_growth_exponent", 2.0),
    ("adaptive_linesearch_parameters.step_size_growth_exponent", NAN),
    ("malitsky_pock_parameters.step_size_downscaling_factor", 0.0), ("malitsky_pock_parameters.step_size_downscaling_factor", 1.0),
    ("malitsky_pock_parameters.step_size_downscaling_factor", NAN), ("malitsky_pock_parameters.step_size_downscaling_factor", 1.0e-300),
    ("malitsky_pock_parameters.linesearch_contraction_factor", 0.0), ("malitsky_pock_parameters.linesearch_contraction_factor", 1.0),
    ("malitsky_pock_parameters.linesearch_contraction_factor", NAN),
    ("malitsky_pock_parameters.step_size_interpolation", -1.0), ("malitsky_pock_parameters.step_size_interpolation", NAN),
    ("malitsky_pock_parameters.step_size_interpolation", 1.0e300),
    ("num_threads", 0), ("verbosity_level", -1), ("log_interval_seconds", -1.0), ("log_interval_seconds", NAN),
    ("major_iteration_frequency", 0), ("termination_check_frequency", 0), ("restart_strategy", 0),
    ("primal_weight_update_smoothing", 1.1), ("primal_weight_update_smoothing", -0.1), ("primal_weight_update_smoothing", NAN),
    ("initial_primal_weight", -1.0), ("initial_primal_weight", NAN), ("initial_primal_weight", 1.0e-300), ("initial_primal_weight", 1.0e300),
    ("l_inf_ruiz_iterations", -1), ("l_inf_ruiz_iterations", 1000),
    ("sufficient_reduction_for_restart", 1.0), ("sufficient_reduction_for_restart", 0.0), ("sufficient_reduction_for_restart", NAN),
    ("necessary_reduction_for_restart", 1.0), ("necessary_reduction_for_restart", NAN),
    ("linesearch_rule", 0),
    ("initial_step_size_scaling", -1.0), ("initial_step_size_scaling", NAN), ("initial_step_size_scaling", 1.0e-300), ("initial_step_size_scaling", 1.0e300),
    ("infinite_constraint_bound_threshold", -1.0), ("infinite_constraint_bound_threshold", NAN),
    ("diagonal_qp_trust_region_solver_tolerance", -1.0), ("diagonal_qp_trust_region_solver_tolerance", NAN),
    ("diagonal_qp_trust_region_solver_tolerance", 1.0e-300),
]


def test_default_is_valid(validate):
    assert validate(P()) == ""


@pytest.mark.parametrize("path,value", BAD, ids=[f"{p.split('.')[-1]}={v}" for p, v in BAD])
def test_bad_value_names_the_field(validate, path, value):
    p = P()
    setter(path, value)(p)
    msg = validate(p)
    assert msg != "" and path.split(".")[-1] in msg, msg


def test_necessary_reduction_below_sufficient(validate):  # :554-560
    p = P()
    p.sufficient_reduction_for_restart = 0.5
    p.necessary_reduction_for_restart = 0.4
    assert "necessary_reduction_for_restart" in validate(p)


@pytest.mark.parametrize("old", ["eps_optimal_absolute", "eps_optimal_relative"])
@pytest.mark.parametrize("new", ["simple_optimality_criteria", "detailed_optimality_criteria"])
def test_deprecated_and_new_criteria_conflict(validate, old, new):  # :139-172
    p = P()
    setattr(p.termination_criteria, old, 1.0)
    getattr(p.termination_criteria, new).SetInParent()
    assert new in validate(p)


def test_feasibility_polishing_options(validate):  # :680-706
    p = P()
    p.use_feasibility_polishing = True
    p.handle_some_primal_gradients_on_finite_bounds_as_residuals = False
    p.presolve_options.use_glop = False
    assert validate(p) == ""
    p.handle_some_primal_gradients_on_finite_bounds_as_residuals = True
    assert "use_feasibility_polishing" in validate(p)
    p.handle_some_primal_gradients_on_finite_bounds_as_residuals = False
    p.presolve_options.use_glop = True
    assert "use_feasibility_polishing" in validate(p)


def test_messages_agree_between_oracle_and_product():
    for path, value in BAD:
        p = P()
        setter(path, value)(p)
        strip = lambda s: s.replace("INVALID_ARGUMENT: ", "")
        assert strip(ob.validate_params(p)) == strip(pdlp_b200.validate_params(p)), path

"""ortools/pdlp/python/pdlp_test.py re-typed against `pdlp_b200` (the module that stands in for
`ortools.pdlp.python.pdlp`, pdlp.cc:39-151): the same attribute assignments (lists where the reference
accepts lists, a scipy CSR matrix for the constraint matrix), the same calls, the same expectations.
The QuadraticProgram / MPModelProto cases are host code and run on CPU; the solves run through the
`backend` fixture (CPU oracle here, the CUDA path with `-m gpu`)."""
import numpy as np
import scipy.sparse

import pdlp_b200 as pdlp
from pdlp_b200 import solve_log_pb2, solvers_pb2


def small_proto_lp():  # pdlp_test.py:26-49: min -2y s.t. x + y <= 1, x, y >= 0
    m = pdlp.MPModelProto()
    m.maximize = False
    m.objective_offset = 0.0
    for name, c in (("x", 0.0), ("y", -2.0)):
        v = m.variable.add()
        v.lower_bound, v.upper_bound, v.objective_coefficient, v.name = 0.0, np.inf, c, name
    con = m.constraint.add()
    con.var_index.extend([0, 1])
    con.coefficient.extend([1.0, 1.0])
    con.lower_bound, con.upper_bound = -np.inf, 1.0
    return m


def small_proto_qp():  # pdlp_test.py:52-76: min 2 x*x s.t. x + y <= 1, x, y >= 0
    m = small_proto_lp()
    m.variable[1].objective_coefficient = 0.0
    m.quadratic_objective.qvar1_index.append(0)
    m.quadratic_objective.qvar2_index.append(0)
    m.quadratic_objective.coefficient.append(2.0)
    return m


def _same_model(a, b):
    """assertEqual on two protos whose defaults were all set explicitly: field by field."""
    assert a.maximize == b.maximize and a.objective_offset == b.objective_offset
    assert len(a.variable) == len(b.variable) and len(a.constraint) == len(b.constraint)
    for va, vb in zip(a.variable, b.variable):
        assert (va.lower_bound, va.upper_bound, va.objective_coefficient, va.name) == (vb.lower_bound, vb.upper_bound, vb.objective_coefficient, vb.name)
    for ca, cb in zip(a.constraint, b.constraint):
        assert list(ca.var_index) == list(cb.var_index) and list(ca.coefficient) == list(cb.coefficient)
        assert (ca.lower_bound, ca.upper_bound) == (cb.lower_bound, cb.upper_bound)
    qa, qb = a.quadratic_objective, b.quadratic_objective
    assert list(qa.qvar1_index) == list(qb.qvar1_index) and list(qa.qvar2_index) == list(qb.qvar2_index)
    assert list(qa.coefficient) == list(qb.coefficient)


# ---- QuadraticProgramTest (pdlp_test.py:78-127) ----

def test_validate_quadratic_program_dimensions_for_empty_qp():
    qp = pdlp.QuadraticProgram()
    qp.resize_and_initialize(3, 2)
    pdlp.validate_quadratic_program_dimensions(qp)
    assert pdlp.is_linear_program(qp)


def test_converts_from_tiny_mpmodel_lp():
    qp = pdlp.qp_from_mpmodel_proto(small_proto_lp(), relax_integer_variables=False)
    pdlp.validate_quadratic_program_dimensions(qp)
    assert pdlp.is_linear_program(qp)
    assert sorted(qp.objective_vector) == [-2, 0]


def test_converts_from_tiny_mpmodel_qp():
    qp = pdlp.qp_from_mpmodel_proto(small_proto_qp(), relax_integer_variables=False)
    pdlp.validate_quadratic_program_dimensions(qp)
    assert not pdlp.is_linear_program(qp)
    assert list(qp.objective_vector) == [0, 0]


def _built(objective_vector):
    qp = pdlp.QuadraticProgram()
    qp.objective_vector = objective_vector
    qp.constraint_matrix = scipy.sparse.csr_matrix(np.array([[1.0, 1.0]]))
    qp.constraint_lower_bounds = [-np.inf]
    qp.constraint_upper_bounds = [1.0]
    qp.variable_lower_bounds = [0.0, 0.0]
    qp.variable_upper_bounds = [np.inf, np.inf]
    qp.variable_names = ["x", "y"]
    return qp


def test_build_lp():
    _same_model(pdlp.qp_to_mpmodel_proto(_built([0, -2])), small_proto_lp())


def test_build_qp():
    qp = _built([0, 0])
    qp.set_objective_matrix_diagonal([4.0])  # shorter than the variable count, as in the reference's test
    _same_model(pdlp.qp_to_mpmodel_proto(qp), small_proto_qp())


# ---- PrimalDualHybridGradientTest (pdlp_test.py:130-248) ----

def tiny_lp():
    qp = pdlp.QuadraticProgram()
    qp.objective_offset = -14
    qp.objective_vector = [5, 2, 1, 1]
    qp.constraint_lower_bounds = [12, 7, 1]
    qp.constraint_upper_bounds = [12, np.inf, np.inf]
    qp.variable_lower_bounds = np.zeros(4)
    qp.variable_upper_bounds = [2, 4, 6, 3]
    qp.constraint_matrix = scipy.sparse.csr_matrix(np.array([[2, 1, 1, 2], [1, 0, 1, 0], [0, 0, 1, -1]]))
    return qp


def the_test_lp():
    qp = pdlp.QuadraticProgram()
    qp.objective_offset = -14
    qp.objective_vector = [5.5, -2, -1, 1]
    qp.constraint_lower_bounds = [12, -np.inf, -4, -1]
    qp.constraint_upper_bounds = [12, 7, np.inf, 1]
    qp.variable_lower_bounds = [-np.inf, -2, -np.inf, 2.5]
    qp.variable_upper_bounds = [np.inf, np.inf, 6, 3.5]
    qp.constraint_matrix = scipy.sparse.csr_matrix(np.array([[2, 1, 1, 2], [1, 0, 1, 0], [4, 0, 0, 0], [0, 0, 1.5, -1]]))
    return qp


def _tight():
    params = solvers_pb2.PrimalDualHybridGradientParams()
    params.termination_criteria.simple_optimality_criteria.eps_optimal_relative = 0.0
    params.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = 1.0e-10
    # Not in the reference's test (its limit is INT32_MAX): a guard that turns a solve that does not converge into a
    # failed assertion instead of a hung suite.  The oracle needs 320 / 192 iterations.
    params.termination_criteria.iteration_limit = 20000
    return params


def _almost(actual, expected):  # assertSequenceAlmostEqual: 7 places
    np.testing.assert_allclose(np.asarray(actual, dtype=float), np.asarray(expected, dtype=float), rtol=0, atol=5e-8)


def test_iteration_limit(backend):
    params = solvers_pb2.PrimalDualHybridGradientParams()
    params.termination_criteria.iteration_limit = 1
    params.termination_check_frequency = 1
    result = backend.solve(tiny_lp(), params)
    assert result.solve_log.iteration_count <= 1
    assert result.solve_log.termination_reason == solve_log_pb2.TERMINATION_REASON_ITERATION_LIMIT


def test_solution(backend):
    result = backend.solve(tiny_lp(), _tight())
    assert result.solve_log.termination_reason == solve_log_pb2.TERMINATION_REASON_OPTIMAL
    _almost(result.primal_solution, [1.0, 0.0, 6.0, 2.0])
    _almost(result.dual_solution, [0.5, 4.0, 0.0])
    _almost(result.reduced_costs, [0.0, 1.5, -3.5, 0.0])


def test_solution_2(backend):
    result = backend.solve(the_test_lp(), _tight())
    assert result.solve_log.termination_reason == solve_log_pb2.TERMINATION_REASON_OPTIMAL
    _almost(result.primal_solution, [-1, 8, 1, 2.5])
    _almost(result.dual_solution, [-2, 0, 2.375, 2 / 3])


def test_starting_point(backend):
    params = _tight()
    params.l_inf_ruiz_iterations = 0
    params.l2_norm_rescaling = False
    result = backend.solve(tiny_lp(), params, initial_solution=([1.0, 0.0, 6.0, 2.0], [0.5, 4.0, 0.0]))
    assert result.solve_log.termination_reason == solve_log_pb2.TERMINATION_REASON_OPTIMAL
    assert result.solve_log.iteration_count == 0


# ---- samples/config1_small_lp.py (BASELINE.json config 1) ----

def test_config1_sample(backend):
    import importlib.util
    import io
    import os

    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "samples", "config1_small_lp.py")
    spec = importlib.util.spec_from_file_location("config1_small_lp", path)
    sample = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(sample)
    text = io.StringIO()
    result = sample.run(solve=backend.solve, out=text)
    lines = text.getvalue().splitlines()
    assert lines[0].startswith("TERMINATION_REASON_OPTIMAL after ") and "returned point: POINT_TYPE_" in lines[0]
    assert lines[1].startswith("objective: primal ") and abs(float(lines[1].split()[2]) + 34.0) < 1e-4
    np.testing.assert_allclose(result.primal_solution, [-1, 8, 1, 2.5], atol=1e-4)  # test_util.h:45-48
    np.testing.assert_allclose(result.dual_solution, [-2, 0, 2.375, 2 / 3], atol=1e-4)
    assert [line.split()[0] for line in lines[4:]] == ["x0", "x1", "x2", "x3", "y0", "y1", "y2", "y3"]


def test_enum_namespaces_behave_like_generated_modules():
    assert solve_log_pb2.TerminationReason.Name(solve_log_pb2.TERMINATION_REASON_OPTIMAL) == "TERMINATION_REASON_OPTIMAL"
    assert solve_log_pb2.TerminationReason.Value("TERMINATION_REASON_ITERATION_LIMIT") == solve_log_pb2.TERMINATION_REASON_ITERATION_LIMIT
    assert solve_log_pb2.PointType.Name(pdlp.PointType.POINT_TYPE_AVERAGE_ITERATE) == "POINT_TYPE_AVERAGE_ITERATE"
    assert "RESTART_CHOICE_NO_RESTART" in solve_log_pb2.RestartChoice.keys()
    assert solvers_pb2.PrimalDualHybridGradientParams.RestartStrategy.Name(solvers_pb2.PrimalDualHybridGradientParams.ADAPTIVE_HEURISTIC) == "ADAPTIVE_HEURISTIC"

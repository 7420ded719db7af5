"""MPModelProto <-> QuadraticProgram and the MPSolver-facing solve entry point
(SURVEY.md 8(f) rank 2): the format every OR-Tools front end emits, in front of the device
solver.

Mirrors
  * `QpFromMpModelProto` / `QpToMpModelProto` / `CanFitInMpModelProto`
    (/root/reference/ortools/pdlp/quadratic_program.cc:92-313), exposed by the reference's
    pybind module as `qp_from_mpmodel_proto` / `qp_to_mpmodel_proto` (python/pdlp.cc:93-116);
  * `PdlpSolveProto` (ortools/linear_solver/proto_solver/pdlp_proto_solver.cc:36-136): request
    -> params, model extraction / validation, status mapping, dual sign flip for maximisation;
  * the part of `ExtractValidMPModelOrPopulateResponseStatus` / `FindErrorInMPModelProto`
    (ortools/linear_solver/model_validator.cc:557-761) that applies to models PDLP accepts
    (variables, linear constraints, quadratic objective, solution hint).

The message classes are wire-compatible with ortools/linear_solver/linear_solver.proto
(package `operations_research`, proto2); the field tables below are checked field by field
against the reference .proto in tests/test_mp_model.py (fixture
tests/golden/linear_solver_schema.json).  Host-side only: no compute happens here, the solve
goes through `primal_dual_hybrid_gradient` (the C ABI).
"""
import math

import numpy as np
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory, text_format

try:
    import scipy.sparse as _sp
except ImportError:  # pragma: no cover
    _sp = None

_PACKAGE = "operations_research"
_FILE = "ortools/linear_solver/linear_solver.proto"
_INF = float("inf")

# name -> [(number, field, label, type, default, packed)]; type is a scalar name, "enum:<name>" or
# "msg:<name>".  Only the messages reachable from MPModelRequest / MPSolutionResponse that PDLP
# reads or writes are typed; general constraints and model deltas are carried as opaque
# messages (their presence is all this path looks at, quadratic_program.cc:95-97).
SCHEMA = {
    "MPVariableProto": [
        (1, "lower_bound", "optional", "double", "-inf", False),
        (2, "upper_bound", "optional", "double", "inf", False),
        (3, "objective_coefficient", "optional", "double", "0", False),
        (4, "is_integer", "optional", "bool", "false", False),
        (5, "name", "optional", "string", "", False),
        (6, "branching_priority", "optional", "int32", "0", False),
    ],
    "MPConstraintProto": [
        (6, "var_index", "repeated", "int32", None, True),
        (7, "coefficient", "repeated", "double", None, True),
        (2, "lower_bound", "optional", "double", "-inf", False),
        (3, "upper_bound", "optional", "double", "inf", False),
        (4, "name", "optional", "string", "", False),
        (5, "is_lazy", "optional", "bool", "false", False),
    ],
    "MPGeneralConstraintProto": [
        (1, "name", "optional", "string", "", False),
    ],
    "MPQuadraticObjective": [
        (1, "qvar1_index", "repeated", "int32", None, False),
        (2, "qvar2_index", "repeated", "int32", None, False),
        (3, "coefficient", "repeated", "double", None, False),
    ],
    "PartialVariableAssignment": [
        (1, "var_index", "repeated", "int32", None, True),
        (2, "var_value", "repeated", "double", None, True),
    ],
    "MPModelProto": [
        (3, "variable", "repeated", "msg:MPVariableProto", None, False),
        (4, "constraint", "repeated", "msg:MPConstraintProto", None, False),
        (7, "general_constraint", "repeated", "msg:MPGeneralConstraintProto", None, False),
        (1, "maximize", "optional", "bool", "false", False),
        (2, "objective_offset", "optional", "double", "0", False),
        (8, "quadratic_objective", "optional", "msg:MPQuadraticObjective", None, False),
        (5, "name", "optional", "string", "", False),
        (6, "solution_hint", "optional", "msg:PartialVariableAssignment", None, False),
    ],
    "MPModelDeltaProto": [
        (1, "baseline_model_file_path", "optional", "string", None, False),
    ],
    "MPModelRequest": [
        (1, "model", "optional", "msg:MPModelProto", None, False),
        (2, "solver_type", "optional", "enum:MPModelRequest.SolverType", "GLOP_LINEAR_PROGRAMMING", False),
        (3, "solver_time_limit_seconds", "optional", "double", None, False),
        (4, "enable_internal_solver_output", "optional", "bool", "false", False),
        (5, "solver_specific_parameters", "optional", "string", None, False),
        (9, "ignore_solver_specific_parameters_failure", "optional", "bool", "false", False),
        (8, "model_delta", "optional", "msg:MPModelDeltaProto", None, False),
        (11, "populate_additional_solutions_up_to", "optional", "int32", "0", False),
    ],
    "MPSolution": [
        (1, "objective_value", "optional", "double", None, False),
        (2, "variable_value", "repeated", "double", None, True),
    ],
    "MPSolveInfo": [
        (1, "solve_wall_time_seconds", "optional", "double", None, False),
        (2, "solve_user_time_seconds", "optional", "double", None, False),
    ],
    "MPSolutionResponse": [
        (1, "status", "optional", "enum:MPSolverResponseStatus", "MPSOLVER_UNKNOWN_STATUS", False),
        (7, "status_str", "optional", "string", None, False),
        (2, "objective_value", "optional", "double", None, False),
        (5, "best_objective_bound", "optional", "double", None, False),
        (3, "variable_value", "repeated", "double", None, True),
        (10, "solve_info", "optional", "msg:MPSolveInfo", None, False),
        (11, "solver_specific_info", "optional", "bytes", None, False),
        (4, "dual_value", "repeated", "double", None, True),
        (6, "reduced_cost", "repeated", "double", None, True),
        (8, "additional_solutions", "repeated", "msg:MPSolution", None, False),
    ],
}

ENUMS = {
    "MPSolverResponseStatus": {
        "MPSOLVER_OPTIMAL": 0x0, "MPSOLVER_FEASIBLE": 0x1, "MPSOLVER_INFEASIBLE": 0x2, "MPSOLVER_UNBOUNDED": 0x3,
        "MPSOLVER_ABNORMAL": 0x4, "MPSOLVER_NOT_SOLVED": 0x6, "MPSOLVER_MODEL_IS_VALID": 0x61,
        "MPSOLVER_CANCELLED_BY_USER": 0x62, "MPSOLVER_UNKNOWN_STATUS": 0x63, "MPSOLVER_MODEL_INVALID": 0x5,
        "MPSOLVER_MODEL_INVALID_SOLUTION_HINT": 0x54, "MPSOLVER_MODEL_INVALID_SOLVER_PARAMETERS": 0x55,
        "MPSOLVER_SOLVER_TYPE_UNAVAILABLE": 0x7, "MPSOLVER_INCOMPATIBLE_OPTIONS": 0x71,
    },
    "MPModelRequest.SolverType": {
        "CLP_LINEAR_PROGRAMMING": 0, "GLOP_LINEAR_PROGRAMMING": 2, "GLPK_LINEAR_PROGRAMMING": 1,
        "GUROBI_LINEAR_PROGRAMMING": 6, "XPRESS_LINEAR_PROGRAMMING": 101, "CPLEX_LINEAR_PROGRAMMING": 10,
        "HIGHS_LINEAR_PROGRAMMING": 15, "SCIP_MIXED_INTEGER_PROGRAMMING": 3, "GLPK_MIXED_INTEGER_PROGRAMMING": 4,
        "CBC_MIXED_INTEGER_PROGRAMMING": 5, "GUROBI_MIXED_INTEGER_PROGRAMMING": 7,
        "XPRESS_MIXED_INTEGER_PROGRAMMING": 102, "CPLEX_MIXED_INTEGER_PROGRAMMING": 11,
        "HIGHS_MIXED_INTEGER_PROGRAMMING": 16, "BOP_INTEGER_PROGRAMMING": 12, "SAT_INTEGER_PROGRAMMING": 14,
        "PDLP_LINEAR_PROGRAMMING": 8, "KNAPSACK_MIXED_INTEGER_PROGRAMMING": 13,
    },
}

_SCALAR = {
    "double": descriptor_pb2.FieldDescriptorProto.TYPE_DOUBLE,
    "int32": descriptor_pb2.FieldDescriptorProto.TYPE_INT32,
    "bool": descriptor_pb2.FieldDescriptorProto.TYPE_BOOL,
    "string": descriptor_pb2.FieldDescriptorProto.TYPE_STRING,
    "bytes": descriptor_pb2.FieldDescriptorProto.TYPE_BYTES,
}


def _build():
    F = descriptor_pb2.FieldDescriptorProto
    fp = descriptor_pb2.FileDescriptorProto()
    fp.name, fp.package, fp.syntax = _FILE, _PACKAGE, "proto2"
    shells = {}
    for name in SCHEMA:
        m = fp.message_type.add()
        m.name = name
        shells[name] = m
    for ename, values in ENUMS.items():
        if "." in ename:
            owner, short = ename.split(".")
            e = shells[owner].enum_type.add()
        else:
            short = ename
            e = fp.enum_type.add()
        e.name = short
        for vname, number in values.items():
            v = e.value.add()
            v.name, v.number = vname, number
    for name, fields in SCHEMA.items():
        for number, fname, label, ftype, default, packed in fields:
            f = shells[name].field.add()
            f.name, f.number = fname, number
            f.label = F.LABEL_REPEATED if label == "repeated" else F.LABEL_OPTIONAL
            if ftype.startswith("msg:"):
                f.type, f.type_name = F.TYPE_MESSAGE, "." + _PACKAGE + "." + ftype[4:]
            elif ftype.startswith("enum:"):
                f.type, f.type_name = F.TYPE_ENUM, "." + _PACKAGE + "." + ftype[5:]
            else:
                f.type = _SCALAR[ftype]
            if default is not None and label != "repeated" and not ftype.startswith("msg:"):
                f.default_value = default
            if packed:
                f.options.packed = True
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fp)
    return {name: message_factory.GetMessageClass(pool.FindMessageTypeByName(_PACKAGE + "." + name)) for name in SCHEMA}, fp


_classes, file_descriptor_proto = _build()
MPVariableProto = _classes["MPVariableProto"]
MPConstraintProto = _classes["MPConstraintProto"]
MPQuadraticObjective = _classes["MPQuadraticObjective"]
MPModelProto = _classes["MPModelProto"]
MPModelRequest = _classes["MPModelRequest"]
MPSolutionResponse = _classes["MPSolutionResponse"]
for _name, _number in ENUMS["MPSolverResponseStatus"].items():
    globals()[_name] = _number
PDLP_LINEAR_PROGRAMMING = ENUMS["MPModelRequest.SolverType"]["PDLP_LINEAR_PROGRAMMING"]

_INT32_MAX = 2**31 - 1


def qp_from_mpmodel_proto(proto, relax_integer_variables, include_names=False):
    """quadratic_program.cc:92-206.  Raises ValueError with the reference's message
    (the pybind wrapper turns the absl status into std::invalid_argument, python/pdlp.cc:93-106)."""
    from ._api import QuadraticProgram

    if len(proto.general_constraint) > 0:
        raise ValueError("General constraints are not supported.")
    primal_size = len(proto.variable)
    dual_size = len(proto.constraint)
    qp = QuadraticProgram()
    qp.resize_and_initialize(primal_size, dual_size)
    if include_names:
        qp.problem_name = proto.name
        qp.variable_names = [""] * primal_size
        qp.constraint_names = [""] * dual_size
    for i, var in enumerate(proto.variable):
        qp.variable_lower_bounds[i] = var.lower_bound
        qp.variable_upper_bounds[i] = var.upper_bound
        qp.objective_vector[i] = var.objective_coefficient
        if var.is_integer and not relax_integer_variables:
            raise ValueError("Integer variable encountered with relax_integer_variables == false")
        if include_names:
            qp.variable_names[i] = var.name
    counts = np.zeros(dual_size + 1, dtype=np.int64)
    for i, con in enumerate(proto.constraint):
        for j, v in enumerate(con.var_index):
            if v < 0 or v >= primal_size:
                raise ValueError(f"Variable index of {i}th constraint's {j}th nonzero is {v} which is not in the allowed range [0, {primal_size})")
        qp.constraint_lower_bounds[i] = con.lower_bound
        qp.constraint_upper_bounds[i] = con.upper_bound
        if include_names:
            qp.constraint_names[i] = con.name
        if len(con.var_index) != len(con.coefficient):
            raise ValueError(f"{i}th constraint has {len(con.coefficient)} coefficients, expected {len(con.var_index)}")
        counts[i + 1] = len(con.var_index)
    # Row-compressed assembly then one conversion: the result is the compressed column-major
    # matrix with ascending rows per column that Eigen's insert() + makeCompressed() builds
    # (quadratic_program.cc:134-157).  Repeated (row, column) pairs - rejected by Eigen's
    # insert() with an assertion in debug builds - are summed.
    indptr = np.cumsum(counts)
    indices = np.empty(int(indptr[-1]), dtype=np.int64)
    data = np.empty(int(indptr[-1]), dtype=np.float64)
    for i, con in enumerate(proto.constraint):
        indices[indptr[i]:indptr[i + 1]] = con.var_index
        data[indptr[i]:indptr[i + 1]] = con.coefficient
    csr = _sp.csr_matrix((data, indices, indptr), shape=(dual_size, primal_size))
    csc = csr.tocsc()
    csc.sum_duplicates()
    csc.sort_indices()
    qp.constraint_matrix = csc
    quadratic = proto.quadratic_objective
    n1, n2, nc = len(quadratic.qvar1_index), len(quadratic.qvar2_index), len(quadratic.coefficient)
    if n1 != n2 or n1 != nc:
        raise ValueError(f"The quadratic objective has {n1} qvar1_indices, {n2} qvar2_indices, and {nc} coefficients, expected equal numbers.")
    diagonal = np.zeros(primal_size) if n1 > 0 else None
    for i in range(n1):
        index1, index2 = quadratic.qvar1_index[i], quadratic.qvar2_index[i]
        if index1 < 0 or index2 < 0 or index1 >= primal_size or index2 >= primal_size:
            raise ValueError(f"The quadratic objective's {i}th nonzero has indices {index1} and {index2}, which are not both in the expected range [0, {primal_size})")
        if index1 != index2:
            raise ValueError(f"The quadratic objective's {i}th nonzero has off-diagonal element at ({index1}, {index2}). Only diagonal objective matrices are supported.")
        # QuadraticProgram has an implicit 1/2 in front of the quadratic term.
        diagonal[index1] = 2 * quadratic.coefficient[i]
    if diagonal is not None:
        qp.set_objective_matrix_diagonal(diagonal)
    qp.objective_offset = proto.objective_offset
    if proto.maximize:
        qp.objective_offset *= -1
        qp.objective_vector = -qp.objective_vector
        if diagonal is not None:
            qp.set_objective_matrix_diagonal(-diagonal)
        qp.objective_scaling_factor = -1.0
    return qp


def can_fit_in_mpmodel_proto(qp, largest_ok_size=_INT32_MAX):
    """quadratic_program.cc:208-231 (`internal::TestableCanFitInMpModelProto`)."""
    primal_size, dual_size = len(qp.variable_lower_bounds), len(qp.constraint_lower_bounds)
    if primal_size > largest_ok_size:
        raise ValueError(f"Too many variables ({primal_size}) to index with an int32_t.")
    if dual_size > largest_ok_size:
        raise ValueError(f"Too many constraints ({dual_size}) to index with an int32_t.")


def qp_to_mpmodel_proto(qp):
    """quadratic_program.cc:233-313."""
    can_fit_in_mpmodel_proto(qp)
    if qp.objective_scaling_factor == 0:
        raise ValueError("objective_scaling_factor cannot be zero.")
    s = qp.objective_scaling_factor
    proto = MPModelProto()
    if qp.problem_name:
        proto.name = qp.problem_name
    proto.objective_offset = s * qp.objective_offset
    proto.maximize = s < 0
    for i in range(len(qp.variable_lower_bounds)):
        var = proto.variable.add()
        var.lower_bound = float(qp.variable_lower_bounds[i])
        var.upper_bound = float(qp.variable_upper_bounds[i])
        var.objective_coefficient = float(s * qp.objective_vector[i])
        if qp.variable_names is not None and i < len(qp.variable_names) and qp.variable_names[i]:
            var.name = qp.variable_names[i]
    for i in range(len(qp.constraint_lower_bounds)):
        con = proto.constraint.add()
        con.lower_bound = float(qp.constraint_lower_bounds[i])
        con.upper_bound = float(qp.constraint_upper_bounds[i])
        if qp.constraint_names is not None and i < len(qp.constraint_names) and qp.constraint_names[i]:
            con.name = qp.constraint_names[i]
    # Column by column, as the reference iterates (each constraint's entries end up in ascending
    # variable order).
    csc = _sp.csc_matrix(qp.constraint_matrix)
    csc.sort_indices()
    for col in range(csc.shape[1]):
        for k in range(csc.indptr[col], csc.indptr[col + 1]):
            con = proto.constraint[int(csc.indices[k])]
            con.var_index.append(col)
            con.coefficient.append(float(csc.data[k]))
    if qp.objective_matrix is not None:
        quadratic = proto.quadratic_objective
        quadratic.SetInParent()
        for i, d in enumerate(qp.objective_matrix):
            if d != 0.0:
                quadratic.qvar1_index.append(i)
                quadratic.qvar2_index.append(i)
                quadratic.coefficient.append(float(s * d / 2.0))
    return proto


# ---- model validation (the subset of model_validator.cc a PDLP request can reach) ----
_VALIDATOR_INFINITY = 1e100  # FLAGS_model_validator_infinity, model_validator.cc:42-44


def _bad(value, threshold):
    return math.isnan(value) or abs(value) >= threshold


def _fmt_f(value):
    return "%f" % value


def _bounds_error(element, threshold):
    """FindErrorInBounds, model_validator.cc:55-68."""
    lo, hi = element.lower_bound, element.upper_bound
    if math.isnan(lo) or math.isnan(hi) or lo >= threshold or hi <= -threshold or lo > hi:
        return f"Infeasible bounds: [{_fmt_f(lo)}, {_fmt_f(hi)}]"
    return ""


def _short(message):
    return text_format.MessageToString(message, as_one_line=True)


def find_error_in_mp_model_proto(model, abs_value_threshold=0.0):
    """FindErrorInMPModelProto, model_validator.cc:557-683, for models without general
    constraints and annotations (PDLP rejects the former, quadratic_program.cc:95-97)."""
    threshold = abs_value_threshold or _VALIDATOR_INFINITY
    if _bad(model.objective_offset, threshold):
        return f"Invalid objective_offset: {model.objective_offset:g}"
    num_vars = len(model.variable)
    for i, variable in enumerate(model.variable):
        error = _bounds_error(variable, threshold)
        if not error and variable.is_integer and np.ceil(variable.lower_bound) > np.floor(variable.upper_bound):
            error = f"Infeasible bounds for integer variable: [{variable.lower_bound:g}, {variable.upper_bound:g}] translate to the empty set"
        if not error and _bad(variable.objective_coefficient, threshold):
            error = f"Invalid objective_coefficient: {variable.objective_coefficient:g}"
        if error:
            return f"In variable #{i}: {error}. Variable proto: {_short(variable)}"
    for i, constraint in enumerate(model.constraint):
        error = _bounds_error(constraint, threshold)
        if not error and len(constraint.var_index) != len(constraint.coefficient):
            error = f"var_index_size() != coefficient_size() ({len(constraint.var_index)} VS {len(constraint.coefficient)}"
        if not error:
            seen = set()
            duplicate = -1
            for j, v in enumerate(constraint.var_index):
                if v >= num_vars or v < 0:
                    error = f"var_index({j})={v} is out of bounds"
                    break
                if _bad(constraint.coefficient[j], threshold):
                    error = f"coefficient({j})={constraint.coefficient[j]:g} is invalid"
                    break
                if v in seen:
                    duplicate = v
                seen.add(v)
            if not error and duplicate >= 0:
                error = f"var_index #{duplicate} appears several times"
        if error:
            cropped = MPConstraintProto()
            cropped.CopyFrom(constraint)
            suffix = ""
            if len(constraint.var_index) > 10:
                del cropped.var_index[10:]
                suffix += f" (var_index cropped; size={len(constraint.var_index)})."
            if len(constraint.coefficient) > 10:
                del cropped.coefficient[10:]
                suffix += f" (coefficient cropped; size={len(constraint.coefficient)})."
            return f"In constraint #{i}: {error}. Constraint proto: {_short(cropped)}{suffix}"
    if len(model.general_constraint) > 0:
        return ""  # validated by the caller's rejection (PDLP has no general constraints)
    if model.HasField("quadratic_objective"):
        q = model.quadratic_objective
        if len(q.qvar1_index) != len(q.qvar2_index) or len(q.qvar1_index) != len(q.coefficient):
            return "In quadratic_objective: indices and coefficients must have the same size"
        for i in range(len(q.qvar1_index)):
            v1, v2 = q.qvar1_index[i], q.qvar2_index[i]
            if v1 >= num_vars or v1 < 0:
                return f"In quadratic_objective: qvar1_index({i})={v1} is invalid. It must be in [0, {num_vars})"
            if v2 >= num_vars or v2 < 0:
                return f"In quadratic_objective: qvar2_index({i})={v2} is invalid. It must be in [0, {num_vars})"
            if _bad(q.coefficient[i], threshold):
                return f"In quadratic_objective: coefficient({i})={q.coefficient[i]:g} is invalid"
    hint = model.solution_hint
    if len(hint.var_index) != len(hint.var_value):
        return f"In solution_hint(): var_index_size() != var_value_size() [{len(hint.var_index)} VS {len(hint.var_value)}"
    seen = set()
    for i, v in enumerate(hint.var_index):
        if v >= num_vars or v < 0:
            return f"In solution_hint(): var_index({i})={v} is invalid. It must be in [0, {num_vars})"
        if v in seen:
            return f"In solution_hint(): Duplicate var_index = {v}"
        seen.add(v)
        if _bad(hint.var_value[i], threshold):
            return f"In solution_hint(): var_value({i})={hint.var_value[i]:g} is invalid"
    return ""


def extract_valid_mp_model(request, response):
    """ExtractValidMPModelOrPopulateResponseStatus, model_validator.cc:685-761.  Returns the model,
    or None after filling `response`.  `model_delta` requests are answered MODEL_INVALID: applying a
    delta needs the baseline file reader, which is outside this path."""
    has_model = request.HasField("model")
    has_delta = request.HasField("model_delta")
    if not has_model and not has_delta:
        response.status = MPSOLVER_OPTIMAL  # noqa: F821
        response.status_str = "Requests without model are considered OPTIMAL"
        return None
    if has_model and has_delta:
        response.status = MPSOLVER_MODEL_INVALID  # noqa: F821
        response.status_str = "Fields 'model' and 'model_delta' are mutually exclusive"
        return None
    if has_delta:
        response.status = MPSOLVER_MODEL_INVALID  # noqa: F821
        response.status_str = "Error when reading model_delta.baseline_model_file_path: 'model deltas are not supported by pdlp_b200"
        return None
    model = request.model
    error = find_error_in_mp_model_proto(model)
    if error:
        response.status = MPSOLVER_INFEASIBLE if "Infeasible" in error else MPSOLVER_MODEL_INVALID  # noqa: F821
        response.status_str = error
        return None
    if len(model.variable) == 0 and len(model.constraint) == 0 and len(model.general_constraint) == 0:
        response.status = MPSOLVER_OPTIMAL  # noqa: F821
        response.objective_value = model.objective_offset
        response.best_objective_bound = response.objective_value
        response.status_str = "Requests without variables and constraints are considered OPTIMAL"
        return None
    return model


def pdlp_solve_proto(request, relax_integer_variables=False, interrupt_solve=None, device=-1, solver=None):
    """`PdlpSolveProto`, pdlp_proto_solver.cc:36-136: MPModelRequest -> MPSolutionResponse through
    the device solver.  `interrupt_solve` is what `primal_dual_hybrid_gradient` accepts (a numpy
    int32 array of one element, non-zero = stop) or None.  Conversion failures raise ValueError
    (the reference returns the absl status); everything else is reported in the response.
    `solver(qp, params, interrupt_solve)` replaces the device solve (the tests run the same body
    against the CPU oracle that way); the default is `primal_dual_hybrid_gradient` on `device`."""
    import pdlp_b200 as _pkg
    from . import _api

    params = _pkg.PrimalDualHybridGradientParams()
    params.verbosity_level = 3 if request.enable_internal_solver_output else 0
    response = MPSolutionResponse()
    try:
        text_format.Merge(request.solver_specific_parameters, params)
    except text_format.ParseError:
        response.status = MPSOLVER_MODEL_INVALID_SOLVER_PARAMETERS  # noqa: F821
        return response
    if interrupt_solve is not None and bool(np.asarray(interrupt_solve).ravel()[0]):
        response.status = MPSOLVER_NOT_SOLVED  # noqa: F821
        return response
    if request.HasField("solver_time_limit_seconds"):
        params.termination_criteria.time_sec_limit = request.solver_time_limit_seconds
    model = extract_valid_mp_model(request, response)
    if model is None:
        return response
    qp = qp_from_mpmodel_proto(model, relax_integer_variables)
    objective_scaling_factor = qp.objective_scaling_factor
    if solver is not None:
        result = solver(qp, params, interrupt_solve)
    else:
        result = _api.primal_dual_hybrid_gradient(qp, params, interrupt_solve=interrupt_solve, device=device)
    return response_from_result(result, objective_scaling_factor)


def response_from_result(result, objective_scaling_factor):
    """The status mapping and sign conventions of pdlp_proto_solver.cc:82-133."""
    import pdlp_b200 as _pkg

    reason = _pkg.TerminationReason
    log = result.solve_log
    response = MPSolutionResponse()
    response.status = {
        reason.TERMINATION_REASON_OPTIMAL: MPSOLVER_OPTIMAL,  # noqa: F821
        reason.TERMINATION_REASON_NUMERICAL_ERROR: MPSOLVER_ABNORMAL,  # noqa: F821
        reason.TERMINATION_REASON_PRIMAL_INFEASIBLE: MPSOLVER_INFEASIBLE,  # noqa: F821
        reason.TERMINATION_REASON_INTERRUPTED_BY_USER: MPSOLVER_CANCELLED_BY_USER,  # noqa: F821
    }.get(log.termination_reason, MPSOLVER_NOT_SOLVED)  # noqa: F821
    if log.HasField("termination_string"):
        response.status_str = log.termination_string
    # GetConvergenceInformation(solution_stats, solution_type), iteration_stats.cc:520-532.
    for info in log.solution_stats.convergence_information:
        if info.candidate_type == log.solution_type:
            response.objective_value = info.primal_objective
            break
    # variable_value and dual_value are set in all cases, as the reference does.
    response.variable_value.extend(np.asarray(result.primal_solution, dtype=np.float64).tolist())
    # Maximisation problems were negated on the way in; duals and reduced costs flip back.
    response.dual_value.extend((objective_scaling_factor * np.asarray(result.dual_solution, dtype=np.float64)).tolist())
    response.reduced_cost.extend((objective_scaling_factor * np.asarray(result.reduced_costs, dtype=np.float64)).tolist())
    response.solver_specific_info = log.SerializeToString()
    return response

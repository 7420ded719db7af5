"""pdlp_b200: Python surface of the B200-native PDLP solver.

Mirrors the reference's pybind11 module `ortools.pdlp.python.pdlp`
(/root/reference/ortools/pdlp/python/pdlp.cc:39-151): `QuadraticProgram`,
`PrimalAndDualSolution`, `SolverResult`, `validate_quadratic_program_dimensions`,
`is_linear_program`, `primal_dual_hybrid_gradient(qp, params, initial_solution=None)`,
plus the two protobuf schemas as `solvers_pb2` / `solve_log_pb2` namespaces.

Everything numeric runs in the native CUDA library through its C ABI
(include/pdlp_b200.h).  There is no Python or CPU fallback: importing this package
without the built library raises ImportError, and solving without a CUDA device raises
RuntimeError.
"""
import ctypes
import os
import types

import numpy as np

from . import _schema

_HERE = os.path.dirname(os.path.abspath(__file__))
LIBRARY_PATH = os.environ.get("PDLP_B200_LIBRARY") or os.path.join(_HERE, "lib", "libpdlp_b200.so")  # override: kernel-variant experiments

if not os.path.exists(LIBRARY_PATH):
    raise ImportError(
        f"{LIBRARY_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(or `make -C or-tools_b200/csrc`). pdlp_b200 has no fallback path."
    )
_lib = ctypes.CDLL(LIBRARY_PATH)


def _prefer_bundled_nccl():
    """The library resolves NCCL with dlopen at the first sharded call.  In a Python process torch
    may be imported later and needs ITS NCCL build under the soname libnccl.so.2, so point the
    library at the copy torch ships (the nvidia-nccl wheel) instead of a system-wide one."""
    if os.environ.get("PDLP_B200_NCCL_LIBRARY"):
        return
    try:
        import importlib.util

        spec = importlib.util.find_spec("nvidia.nccl")
        for root in (spec.submodule_search_locations if spec is not None else []):
            path = os.path.join(root, "lib", "libnccl.so.2")
            if os.path.exists(path):
                os.environ["PDLP_B200_NCCL_LIBRARY"] = path
                return
    except (ImportError, ValueError):
        pass


_prefer_bundled_nccl()

_c_double_p = ctypes.POINTER(ctypes.c_double)
_c_int64_p = ctypes.POINTER(ctypes.c_int64)
_c_uint8_p = ctypes.POINTER(ctypes.c_uint8)


class _CProblem(ctypes.Structure):
    _fields_ = [
        ("num_variables", ctypes.c_int64),
        ("num_constraints", ctypes.c_int64),
        ("num_nonzeros", ctypes.c_int64),
        ("col_ptr", _c_int64_p),
        ("row_idx", _c_int64_p),
        ("values", _c_double_p),
        ("objective_vector", _c_double_p),
        ("objective_matrix_diagonal", _c_double_p),
        ("variable_lower_bounds", _c_double_p),
        ("variable_upper_bounds", _c_double_p),
        ("constraint_lower_bounds", _c_double_p),
        ("constraint_upper_bounds", _c_double_p),
        ("objective_offset", ctypes.c_double),
        ("objective_scaling_factor", ctypes.c_double),
        ("problem_name", ctypes.c_char_p),
        ("col_ptr32", ctypes.POINTER(ctypes.c_int32)),
        ("row_idx32", ctypes.POINTER(ctypes.c_int32)),
    ]


class _CCounters(ctypes.Structure):
    _fields_ = [
        ("kernel_launches", ctypes.c_longlong),
        ("trial_steps", ctypes.c_longlong),
        ("setup_seconds", ctypes.c_double),
        ("loop_seconds", ctypes.c_double),
        ("device_loop_seconds", ctypes.c_double),
        ("collectives", ctypes.c_longlong),
        ("upload_seconds", ctypes.c_double),
        ("symmetric_buffers", ctypes.c_int),
        ("peer_exchanges", ctypes.c_longlong),
        ("interrupts_seen_on_device", ctypes.c_int),
    ]


UNIQUE_ID_BYTES = 128


class _CShard(ctypes.Structure):
    _fields_ = [
        ("rank", ctypes.c_int),
        ("world", ctypes.c_int),
        ("global_num_variables", ctypes.c_int64),
        ("column_offset", ctypes.c_int64),
        ("nccl_unique_id", ctypes.c_void_p),
        ("num_shared_constraints", ctypes.c_int64),
    ]


CALLBACK_TYPE = ctypes.CFUNCTYPE(None, ctypes.c_int, _c_uint8_p, ctypes.c_size_t, _c_double_p, ctypes.c_void_p)

_lib.pdlp_b200_last_error.restype = ctypes.c_char_p
_lib.pdlp_b200_free.argtypes = [ctypes.c_void_p]
_lib.pdlp_b200_free.restype = None
_lib.pdlp_b200_describe_schema.argtypes = [ctypes.POINTER(ctypes.c_void_p)]
_lib.pdlp_b200_solve.argtypes = [
    ctypes.POINTER(_CProblem), ctypes.c_char_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
    ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
    ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_size_t), ctypes.POINTER(_CCounters),
]
_lib.pdlp_b200_solve_sharded.argtypes = [ctypes.POINTER(_CProblem), ctypes.POINTER(_CShard)] + _lib.pdlp_b200_solve.argtypes[1:]
# (original, presolved or NULL, presolve_status, postsolve, postsolve_user) + the tail of pdlp_b200_solve.
_lib.pdlp_b200_solve_presolved.argtypes = [ctypes.POINTER(_CProblem), ctypes.POINTER(_CProblem), ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p] + _lib.pdlp_b200_solve.argtypes[1:]
_lib.pdlp_b200_compute_statuses.argtypes = [ctypes.POINTER(_CProblem), ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
_lib.pdlp_b200_problem_create_sharded.argtypes = [ctypes.POINTER(_CProblem), ctypes.POINTER(_CShard), ctypes.c_int, ctypes.POINTER(ctypes.c_void_p)]
_lib.pdlp_b200_unique_id.argtypes = [ctypes.c_void_p]
_lib.pdlp_b200_validate_params.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_void_p)]
_lib.pdlp_b200_default_params.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_size_t)]

# Device pointers in, device pointer out, nothing allocated (include/pdlp_b200.h); NO_ALLOC_MAX_ROWS = PDLP_B200_NO_ALLOC_MAX_ROWS.
_lib.pdlp_b200_matvec_no_alloc.argtypes = [ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                           ctypes.c_int, ctypes.c_void_p]
NO_ALLOC_MAX_ROWS = 262144

EXPORTED_SYMBOLS = [
    "pdlp_b200_solve", "pdlp_b200_validate_params", "pdlp_b200_default_params", "pdlp_b200_describe_schema",
    "pdlp_b200_last_error", "pdlp_b200_free", "pdlp_b200_problem_create", "pdlp_b200_problem_destroy",
    "pdlp_b200_matvec", "pdlp_b200_scaled_norms", "pdlp_b200_stats", "pdlp_b200_rescale", "pdlp_b200_kkt_stats",
    "pdlp_b200_reduced_costs", "pdlp_b200_localized_bounds", "pdlp_b200_trial_steps", "pdlp_b200_trial_steps_profiled",
    "pdlp_b200_check_termination", "pdlp_b200_unique_id", "pdlp_b200_solve_sharded", "pdlp_b200_problem_create_sharded", "pdlp_b200_release_communicators",
    "pdlp_b200_work_blocks", "pdlp_b200_matvec_no_alloc", "pdlp_b200_format_iteration_stats", "pdlp_b200_localized_bounds_max_norm", "pdlp_b200_solve_presolved",
    "pdlp_b200_compute_statuses",
]


def _check(status):
    if status != 0:
        msg = _lib.pdlp_b200_last_error().decode("utf-8", "replace")
        if status == 1:
            raise ValueError(msg)
        raise RuntimeError(f"pdlp_b200 status {status}: {msg}")


def _take_bytes(ptr, length):
    data = ctypes.string_at(ptr, length) if length else b""
    _lib.pdlp_b200_free(ptr)
    return data


def _describe():
    p = ctypes.c_void_p()
    _check(_lib.pdlp_b200_describe_schema(ctypes.byref(p)))
    text = ctypes.string_at(p).decode()
    _lib.pdlp_b200_free(p)
    return text


_classes, _enums, _file_protos, _enum_wrappers = _schema.build(_describe())


def _namespace(file_key, names):
    ns = types.SimpleNamespace()
    for n in names:
        if "." not in n:
            setattr(ns, n, _classes[n])
    for ename, values in _enums.items():
        if "." in ename:
            continue
        setattr(ns, ename, _enum_wrappers[ename])  # solve_log_pb2.TerminationReason.Name(...) as in a generated module
        for k, v in values.items():
            setattr(ns, k, v)
    return ns


solvers_pb2 = _namespace("solvers", [n for n in _classes])
solve_log_pb2 = solvers_pb2  # one pool; both module names resolve to the same classes
PrimalDualHybridGradientParams = _classes["PrimalDualHybridGradientParams"]
TerminationCriteria = _classes["TerminationCriteria"]
SolveLog = _classes["SolveLog"]
IterationStats = _classes["IterationStats"]
QuadraticProgramStats = _classes["QuadraticProgramStats"]
TerminationReason = _enum_wrappers["TerminationReason"]
PointType = _enum_wrappers["PointType"]
RestartChoice = _enum_wrappers["RestartChoice"]

from ._api import (  # noqa: E402
    ConstraintStatus,
    DeviceProblem,
    GlopProblemStatus,
    glop_status_to_termination_reason,
    VariableStatus,
    compute_statuses,
    IterationCallbackInfo,
    PrimalAndDualSolution,
    QuadraticProgram,
    SolverResult,
    check_termination,
    work_blocks,
    is_linear_program,
    last_solve_counters,
    primal_dual_hybrid_gradient,
    primal_dual_hybrid_gradient_presolved,
    validate_params,
    validate_quadratic_program_dimensions,
)
from . import mp_model  # noqa: E402
from .mp_model import (  # noqa: E402
    MPModelProto,
    MPModelRequest,
    MPSolutionResponse,
    pdlp_solve_proto,
    qp_from_mpmodel_proto,
    qp_to_mpmodel_proto,
)
from . import qp_io  # noqa: E402
from .qp_io import combine_repeated_triplets_in_place, read_quadratic_program_or_die, set_matrix_from_triplets  # noqa: E402
from .sharding import (  # noqa: E402
    ColumnShard,
    RowSharing,
    column_partition,
    primal_dual_hybrid_gradient_sharded,
    release_communicators,
    take_columns,
    unique_id,
)

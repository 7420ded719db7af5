"""CPU-side checks of the drop-in boundary: the C-ABI library loads without a GPU, exports
every symbol include/pdlp_b200.h declares, speaks the reference's proto2 wire format for
PrimalDualHybridGradientParams / SolveLog (tests/golden/pdlp_schema.json, generated from the
reference's .proto files by tests/golden/make_schema_golden.py), and fails loudly -- no CPU
fallback -- when asked to compute without a CUDA device."""
import ctypes
import json
import math
import os
import re

import numpy as np
import pytest

import fixtures
import pdlp_b200

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P = pdlp_b200.PrimalDualHybridGradientParams


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "pdlp_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pdlp_b200_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(pdlp_b200.LIBRARY_PATH)
    names = declared_symbols()
    assert len(names) >= 17
    for name in names:
        assert hasattr(lib, name), name
    assert sorted(pdlp_b200.EXPORTED_SYMBOLS) == [n for n in names if n != "pdlp_b200_callback"]


def test_library_has_no_unresolved_cuda_runtime_dependency():
    # static cudart: the library must load on a box without libcudart.so
    out = os.popen(f"ldd {pdlp_b200.LIBRARY_PATH}").read()
    assert "libcudart" not in out and "not found" not in out, out


def parse_description(text):
    enums, messages = {}, {}
    cur, kind = None, None
    for line in text.splitlines():
        parts = line.split(" ")
        if parts[0] in ("enum", "message"):
            kind, cur = parts[0], parts[1]
            (enums if kind == "enum" else messages)[cur] = {} if kind == "enum" else []
        elif parts[0] == "value":
            enums[cur][parts[1]] = int(parts[2])
        elif parts[0] == "field":
            _, number, name, label, typ, type_name, default, packed, oneof = parts
            messages[cur].append(dict(number=int(number), name=name, label=label, kind=typ, type_name=None if type_name == "-" else type_name,
                                      default=None if default == "-" else default, packed=packed == "1", oneof=None if oneof == "-" else oneof))
    return enums, messages


def test_schema_matches_reference_protos():
    golden = json.load(open(os.path.join(ROOT, "tests", "golden", "pdlp_schema.json")))
    enums, messages = parse_description(pdlp_b200._describe())
    assert enums == golden["enums"]
    assert sorted(messages) == sorted(golden["messages"])
    for name, want_fields in golden["messages"].items():
        got = {f["number"]: f for f in messages[name]}
        assert sorted(got) == [f["number"] for f in want_fields], name
        for want in want_fields:
            g = got[want["number"]]
            for key in ("name", "label", "kind", "type_name", "packed", "oneof"):
                assert g[key] == want[key], (name, want["name"], key, g[key], want[key])
            if want["kind"] == "bool" and want["default"] == "false":
                assert g["default"] in (None, "false", "0"), (name, want["name"])  # an explicit [default = false] is proto2's implicit default
            elif want["kind"] in ("int32", "int64") and want["default"] in (None, "0"):
                assert g["default"] in (None, "0"), (name, want["name"])  # likewise [default = 0]
            elif want["kind"] in ("double", "float") and (want["default"] is None or float(want["default"]) == 0.0):
                assert g["default"] is None or float(g["default"]) == 0.0, (name, want["name"])
            elif want["default"] is None:
                assert g["default"] is None, (name, want["name"])
            elif want["kind"] in ("double", "float"):
                assert float(g["default"]) == float(want["default"]), (name, want["name"])
            elif want["kind"] == "bool":
                assert g["default"] in (want["default"], {"true": "1", "false": "0"}[want["default"]]), (name, want["name"])
            else:
                assert g["default"] == want["default"], (name, want["name"])


def test_default_params_round_trip_through_the_native_codec():
    ptr, ln = ctypes.c_void_p(), ctypes.c_size_t()
    assert pdlp_b200._lib.pdlp_b200_default_params(ctypes.byref(ptr), ctypes.byref(ln)) == 0
    data = pdlp_b200._take_bytes(ptr, ln.value)
    p = P()
    p.ParseFromString(data)
    # the native defaults are the proto defaults (solvers.proto:105-480)
    assert p.termination_criteria.eps_primal_infeasible == 1e-8 and p.major_iteration_frequency == 64 and p.l_inf_ruiz_iterations == 5
    assert p.restart_strategy == P.ADAPTIVE_HEURISTIC and p.linesearch_rule == P.ADAPTIVE_LINESEARCH_RULE
    assert p.adaptive_linesearch_parameters.step_size_reduction_exponent == 0.3 and p.malitsky_pock_parameters.linesearch_contraction_factor == 0.99
    assert math.isinf(p.termination_criteria.time_sec_limit) and p.termination_criteria.iteration_limit == 2**31 - 1


def test_params_with_unknown_and_repeated_fields_parse():
    p = P()
    p.random_projection_seeds.extend([1, 2, 300])
    p.termination_criteria.detailed_optimality_criteria.eps_optimal_objective_gap_absolute = math.inf
    p.num_shards = 7
    # field 1000 (varint) is unknown to the schema: proto2 parsers must skip it
    blob = p.SerializeToString() + bytes([0xC0, 0x3E, 0x05])
    assert pdlp_b200.validate_params(blob) == ""


def test_malformed_params_are_rejected():
    with pytest.raises((ValueError, RuntimeError)):
        pdlp_b200.validate_params(bytes([0x0A, 0xFF, 0xFF]))  # length-delimited field running past the end


def test_no_cpu_fallback_without_a_cuda_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    p = P()
    p.termination_criteria.iteration_limit = 4
    with pytest.raises(RuntimeError):
        pdlp_b200.primal_dual_hybrid_gradient(fixtures.make_test_lp(), p)
    with pytest.raises(RuntimeError):
        pdlp_b200.DeviceProblem(fixtures.make_test_lp())


def test_invalid_inputs_are_reported_before_touching_the_device():
    # these return a SolverResult with the reference's termination reason even with no GPU
    TR = pdlp_b200.TerminationReason
    p = P()
    p.num_threads = 0
    assert pdlp_b200.primal_dual_hybrid_gradient(fixtures.tiny_lp(), p).solve_log.termination_reason == TR.TERMINATION_REASON_INVALID_PARAMETER
    qp = fixtures.tiny_lp()
    qp.objective_vector = np.zeros(0)
    assert pdlp_b200.primal_dual_hybrid_gradient(qp, P()).solve_log.termination_reason == TR.TERMINATION_REASON_INVALID_PROBLEM
    init =pdlp_b200.PrimalAndDualSolution(np.ones(3), np.ones(4))
    out = pdlp_b200.primal_dual_hybrid_gradient(fixtures.make_test_lp(), P(), initial_solution=init)
    assert out.solve_log.termination_reason == TR.TERMINATION_REASON_INVALID_INITIAL_SOLUTION


def test_python_surface_mirrors_the_reference_module():
    # ortools/pdlp/python/pdlp.cc:48-150
    for name in ("QuadraticProgram", "PrimalAndDualSolution", "SolverResult", "primal_dual_hybrid_gradient", "validate_quadratic_program_dimensions",
                 "is_linear_program"):
        assert hasattr(pdlp_b200, name), name
    qp = fixtures.diagonal_qp1()
    assert not pdlp_b200.is_linear_program(qp)
    qp.clear_objective_matrix()
    assert pdlp_b200.is_linear_program(qp)
    pdlp_b200.validate_quadratic_program_dimensions(qp)
    qp.objective_vector = np.zeros(3)
    with pytest.raises(ValueError):
        pdlp_b200.validate_quadratic_program_dimensions(qp)
    assert qp.apply_objective_scaling_and_offset(2.0) == qp.objective_scaling_factor * (2.0 + qp.objective_offset)


# ---- quadratic_program_test.cc:63-132: ValidateQuadraticProgramDimensions ----

def _resized(field, size):
    def mutate(qp):
        setattr(qp, field, np.zeros(size))
    return mutate


def _matrix(rows, cols):
    def mutate(qp):
        import scipy.sparse as sp
        qp.constraint_matrix = sp.csc_matrix((rows, cols))
    return mutate


def test_validate_dimensions_accepts_a_valid_problem():  # :63-67
    pdlp_b200.validate_quadratic_program_dimensions(fixtures.diagonal_qp1())


@pytest.mark.parametrize("mutate", [
    _resized("constraint_lower_bounds", 10),   # :69
    _resized("constraint_upper_bounds", 10),   # :77
    _resized("objective_vector", 10),          # :85
    _resized("variable_lower_bounds", 10),     # :93
    _resized("variable_upper_bounds", 10),     # :101
    _matrix(10, 2),                            # :109
    _matrix(2, 10),                            # :117
    lambda qp: qp.set_objective_matrix_diagonal(np.ones(10)),  # :125
], ids=["constraint_lb", "constraint_ub", "objective", "variable_lb", "variable_ub", "matrix_rows", "matrix_cols", "objective_matrix"])
def test_validate_dimensions_rejects_inconsistent_sizes(mutate):
    qp = pdlp_b200.QuadraticProgram()
    qp.resize_and_initialize(2, 3)
    pdlp_b200.validate_quadratic_program_dimensions(qp)
    mutate(qp)
    with pytest.raises(ValueError, match="Inconsistent dimensions"):
        pdlp_b200.validate_quadratic_program_dimensions(qp)


# ---- the header as plain C: what a cgo / JNI / FFI binding compiles against ----

def build_c_caller():
    import subprocess

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    libdir = os.path.join(root, "or-tools_b200", "lib")
    out_dir = os.path.join(root, "tests", "_build")
    os.makedirs(out_dir, exist_ok=True)
    exe, src = os.path.join(out_dir, "c_caller"), os.path.join(root, "tests", "c", "caller.c")
    deps = [src, os.path.join(root, "include", "pdlp_b200.h"), os.path.join(libdir, "libpdlp_b200.so")]
    if not os.path.exists(exe) or os.path.getmtime(exe) < max(os.path.getmtime(d) for d in deps):
        subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I" + os.path.join(root, "include"), src, "-L" + libdir, "-lpdlp_b200", "-lm",
                        "-Wl,-rpath," + libdir, "-o", exe], check=True)
    return exe


def test_header_compiles_as_c99_and_host_entry_points_work_from_c():
    import subprocess

    run = subprocess.run([build_c_caller(), "cpu"], capture_output=True, text=True, timeout=120)
    assert run.returncode == 0 and "cpu ok" in run.stdout, (run.returncode, run.stdout, run.stderr)

"""The iteration log table of verbosity_level >= 2, against lines typed out from the reference's format
strings (primal_dual_hybrid_gradient.cc:98-182: "%#12.6g ..." / "%#10.4g ...", :193-287: "%6d %8.1f %6.1f",
header labels), through the host-only C-ABI helper pdlp_b200_format_iteration_stats.  A GPU test checks that
a solve with verbosity_level 2 writes header + lines + the termination summary to stderr."""
import ctypes
import os
import subprocess
import sys

import pytest

import pdlp_b200

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PT = pdlp_b200.PointType


def fmt(params, stats, norms=(10.0, 20.0, 3.0, 4.0), iteration_type=0, preferred=PT.POINT_TYPE_AVERAGE_ITERATE):
    lib = pdlp_b200._lib
    pb, sb = params.SerializeToString(), stats.SerializeToString()
    bn = (ctypes.c_double * 4)(*norms)
    h, l = ctypes.c_char_p(), ctypes.c_char_p()
    lib.pdlp_b200_format_iteration_stats.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                                     ctypes.POINTER(ctypes.c_char_p), ctypes.POINTER(ctypes.c_char_p)]
    assert lib.pdlp_b200_format_iteration_stats(pb, len(pb), sb, len(sb), bn, iteration_type, preferred, ctypes.byref(h), ctypes.byref(l)) == 0
    return h.value.decode(), l.value.decode()


def sample_stats():
    s = pdlp_b200.IterationStats()
    s.iteration_number = 128
    s.cumulative_kkt_matrix_passes = 131.0
    s.cumulative_time_sec = 2.34
    for kind, obj in ((PT.POINT_TYPE_CURRENT_ITERATE, -33.0), (PT.POINT_TYPE_AVERAGE_ITERATE, -34.5)):
        c = s.convergence_information.add()
        c.candidate_type = kind
        c.primal_objective = obj
        c.dual_objective = obj - 0.25
        c.l2_primal_residual = 0.5
        c.l2_dual_residual = 0.125
        c.l_inf_primal_residual = 0.25
        c.l_inf_dual_residual = 0.0625
        c.l2_primal_variable = 8.0
        c.l2_dual_variable = 3.0
    return s


def params(level, polishing=False):
    p = pdlp_b200.PrimalDualHybridGradientParams()
    p.verbosity_level = level
    p.use_feasibility_polishing = polishing
    soc = p.termination_criteria.simple_optimality_criteria
    soc.eps_optimal_absolute = 1e-4
    soc.eps_optimal_relative = 1e-4
    return p


def test_level_two_table():
    header, line = fmt(params(2), sample_stats())
    assert header == " iter#   time |  rel_p_res  rel_d_res    rel_gap |   prim_obj   dual_obj"
    # relative residuals (termination.cc:239-271 with eps_abs == eps_rel): residual / (1 + norm), gap / (1 + |p| + |d|)
    rel_p, rel_d, gap = 0.5 / (1 + 20.0), 0.125 / (1 + 10.0), 0.25 / (1 + 34.5 + 34.75)
    assert line == "   128    2.3 | " + "%#10.4g %#10.4g %#10.4g | %#10.4g %#10.4g" % (rel_p, rel_d, gap, -34.5, -34.75)


def test_level_three_and_four_columns_and_tags():
    s = sample_stats()
    header, line = fmt(params(3), s)
    assert header == (" iter# kkt_pass   time | rel_prim_res rel_dual_res      rel_gap |   prim_resid   dual_resid      obj_gap |     prim_obj     dual_obj |"
                      "  prim_var_l2  dual_var_l2")
    rel_p, rel_d, gap = 0.5 / 21.0, 0.125 / 11.0, 0.25 / (1 + 34.5 + 34.75)
    assert line == "   128    131.0    2.3 | " + "%#12.6g %#12.6g %#12.6g | %#12.6g %#12.6g %#12.6g | %#12.6g %#12.6g | %#12.6g %#12.6g" % (
        rel_p, rel_d, gap, 0.5, 0.125, 0.25, -34.5, -34.75, 8.0, 3.0)
    header4, line4 = fmt(params(4), s)
    assert header4.startswith("I  iter# kkt_pass") and line4.startswith("A    128    131.0")
    _, cur = fmt(params(4), s, preferred=PT.POINT_TYPE_CURRENT_ITERATE)
    assert cur.startswith("C    128") and "-33.0000" in cur


def test_polishing_phase_column_and_missing_convergence_information():
    s = sample_stats()
    header, line = fmt(params(2, polishing=True), s, iteration_type=1)
    assert header.startswith("f  iter#") and line.startswith("p    128")
    assert fmt(params(2, polishing=True), s, iteration_type=4)[1].startswith("t ")
    s.ClearField("convergence_information")
    assert fmt(params(2), s)[1] == "   128    2.3"
    assert fmt(params(4), s)[1] == "?    128    131.0    2.3"
    # the preferred candidate is missing: the first entry is shown instead (:207-211)
    s2 = sample_stats()
    del s2.convergence_information[1]
    assert "-33.00" in fmt(params(2), s2)[1]


def test_l_inf_norm_selects_the_l_inf_residuals():
    p = params(3)
    p.termination_criteria.optimality_norm = pdlp_b200.solvers_pb2.OPTIMALITY_NORM_L_INF
    _, line = fmt(p, sample_stats())
    assert "%#12.6g %#12.6g" % (0.25 / (1 + 4.0), 0.0625 / (1 + 3.0)) in line and "%#12.6g %#12.6g" % (0.25, 0.0625) in line


@pytest.mark.gpu
def test_a_solve_logs_the_table_and_the_summary_to_stderr():
    code = ("import sys; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
            "import fixtures, pdlp_b200\n"
            "p = pdlp_b200.PrimalDualHybridGradientParams(); p.verbosity_level = 2; p.major_iteration_frequency = 32; p.termination_check_frequency = 32\n"
            "r = pdlp_b200.primal_dual_hybrid_gradient(fixtures.make_test_lp(), p)\n"
            "print('reason', r.solve_log.termination_reason)\n") % (ROOT, os.path.join(ROOT, "tests"))
    run = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert run.returncode == 0 and "reason 1" in run.stdout, run.stderr
    err = run.stderr.splitlines()
    assert "Problem stats before rescaling:" in err and "Problem stats after rescaling:" in err
    assert any(l.startswith("There are 4 variables, 4 constraints, and 9 constraint matrix nonzeros.") for l in err)
    header = " iter#   time |  rel_p_res  rel_d_res    rel_gap |   prim_obj   dual_obj"
    assert err.count(header) >= 2  # the first table line and the final summary
    assert "Termination reason: TERMINATION_REASON_OPTIMAL" in err and "Solution point type: POINT_TYPE_AVERAGE_ITERATE" in err or "Solution point type: POINT_TYPE_CURRENT_ITERATE" in err
    assert "Final solution stats:" in err
    assert any(l.startswith("     0 ") for l in err) and any(l.startswith("    32 ") for l in err)

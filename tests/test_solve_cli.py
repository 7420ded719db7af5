"""`python -m pdlp_b200.solve`, the mirror of the reference's examples/cpp/pdlp_solve.cc:40-135.  Without a
GPU: flag handling, the params default (verbosity_level 2, overridable), the SolveLog writers and the .sol
text.  On the GPU box: TestLp from an MPS file end to end."""
import os
import subprocess
import sys
import types

import numpy as np
import pytest
from google.protobuf import json_format, text_format

import fixtures
import pdlp_b200
from pdlp_b200 import qp_io, solve as cli

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_params_default_to_verbosity_two_and_merge_text_format():
    assert cli.parse_params("").verbosity_level == 2  # pdlp_solve.cc:86-88
    p = cli.parse_params("verbosity_level: 0 termination_criteria { iteration_limit: 7 simple_optimality_criteria { eps_optimal_absolute: 1e-3 } }")
    assert p.verbosity_level == 0 and p.termination_criteria.iteration_limit == 7
    assert p.termination_criteria.simple_optimality_criteria.eps_optimal_absolute == 1e-3
    with pytest.raises(SystemExit, match="Error parsing --params"):
        cli.parse_params("no_such_field: 1")


def test_input_is_required():
    with pytest.raises(SystemExit, match="--input is required"):
        cli.solve("")


def test_solve_log_is_written_in_the_three_formats(tmp_path):
    log = pdlp_b200.SolveLog()
    log.termination_reason = pdlp_b200.TerminationReason.TERMINATION_REASON_OPTIMAL
    log.iteration_count = 128
    log.solve_time_sec = 0.25
    for ext in (".textproto", ".pb", ".json"):
        f = str(tmp_path / ("log" + ext))
        cli.write_solve_log(f, log)
        back = pdlp_b200.SolveLog()
        data = open(f, "rb").read()
        if ext == ".pb":
            back.ParseFromString(data)
        elif ext == ".json":
            json_format.Parse(data.decode(), back)
        else:
            text_format.Parse(data.decode(), back)
        assert back == log
    with pytest.raises(SystemExit, match="Unrecognized file extension"):
        cli.write_solve_log(str(tmp_path / "log.txt"), log)


def test_sol_text_follows_the_miplib_format():
    log = pdlp_b200.SolveLog()
    log.solution_type = pdlp_b200.PointType.POINT_TYPE_AVERAGE_ITERATE
    c = log.solution_stats.convergence_information.add()
    c.candidate_type = pdlp_b200.PointType.POINT_TYPE_AVERAGE_ITERATE
    c.primal_objective = -34.0
    result = types.SimpleNamespace(solve_log=log, primal_solution=np.array([-1.0, 8.0, 1.0, 2.5]))
    qp = fixtures.make_test_lp()
    assert cli.sol_string(qp, result) == "=obj= -34\nvar0 -1\nvar1 8\nvar2 1\nvar3 2.5\n"
    qp.variable_names = ["a", "b", "c", "d"]
    assert cli.sol_string(qp, result).splitlines()[1:] == ["a -1", "b 8", "c 1", "d 2.5"]
    log.solution_stats.ClearField("convergence_information")
    assert cli.sol_string(qp, result) is None  # pdlp_solve.cc:112: nothing is written without convergence information


@pytest.mark.gpu
def test_command_line_solves_test_lp_from_an_mps_file(tmp_path):
    mps = str(tmp_path / "test_lp.mps")
    qp_io.write_linear_program_to_mps(fixtures.make_test_lp(), mps)
    log_file, sol_file = str(tmp_path / "log.json"), str(tmp_path / "x.sol")
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    run = subprocess.run([sys.executable, "-m", "pdlp_b200.solve", "--input", mps, "--solve_log_file", log_file, "--sol_file", sol_file,
                          "--params", "verbosity_level: 0 termination_criteria { simple_optimality_criteria { eps_optimal_absolute: 1e-8 eps_optimal_relative: 1e-8 } }"],
                         capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert run.returncode == 0, run.stderr
    log = pdlp_b200.SolveLog()
    json_format.Parse(open(log_file).read(), log)
    assert log.termination_reason == pdlp_b200.TerminationReason.TERMINATION_REASON_OPTIMAL
    lines = open(sol_file).read().splitlines()
    assert lines[0].startswith("=obj= ") and abs(float(lines[0].split()[1]) + 34.0) < 1e-5
    values = [float(l.split()[1]) for l in lines[1:]]
    np.testing.assert_allclose(values, [-1, 8, 1, 2.5], atol=1e-5)

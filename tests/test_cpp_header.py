"""The public C++ header (include/pdlp_b200/primal_dual_hybrid_gradient.h) as a caller of the reference
sees it: tests/cpp/caller.cc uses only `operations_research::pdlp::` names
(ortools/pdlp/primal_dual_hybrid_gradient.h:140-156, quadratic_program.h:53-143), is compiled with
g++ against include/ and linked to the built library.  Without a GPU the errors that the entry point
finds before touching the device must come back as termination reasons; on the GPU box the same
binary solves TestLp."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "tests", "_build")
LIBDIR = os.path.join(ROOT, "or-tools_b200", "lib")


def build(name="cpp_caller", source=("tests", "cpp", "caller.cc")):
    os.makedirs(BUILD, exist_ok=True)
    exe = os.path.join(BUILD, name)
    src = os.path.join(ROOT, *source)
    deps = [src, os.path.join(LIBDIR, "libpdlp_b200.so")] + [os.path.join(ROOT, "include", "pdlp_b200", h) for h in os.listdir(os.path.join(ROOT, "include", "pdlp_b200"))]
    if not os.path.exists(exe) or os.path.getmtime(exe) < max(os.path.getmtime(d) for d in deps):
        subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", "-I" + os.path.join(ROOT, "include"), src, "-L" + LIBDIR, "-lpdlp_b200",
                        "-Wl,-rpath," + LIBDIR, "-o", exe], check=True)
    return exe


def test_reference_style_caller_compiles_and_reports_errors_without_a_device():
    run = subprocess.run([build(), "cpu"], capture_output=True, text=True, timeout=120)
    assert run.returncode == 0, (run.returncode, run.stdout, run.stderr)
    assert "cpu ok" in run.stdout


@pytest.mark.gpu
def test_reference_style_caller_solves_test_lp_on_the_device():
    run = subprocess.run([build(), "gpu"], capture_output=True, text=True, timeout=300)
    assert run.returncode == 0, (run.returncode, run.stdout, run.stderr)
    assert "gpu ok: OPTIMAL" in run.stdout


def test_the_sample_program_compiles_against_the_header():
    """samples/config1_small_lp.cc: BASELINE.json config 1 (the reference's sample LP) on this header."""
    assert os.path.exists(build("config1_small_lp", ("samples", "config1_small_lp.cc")))


@pytest.mark.gpu
def test_the_sample_program_solves_config_1_on_the_device():
    run = subprocess.run([build("config1_small_lp", ("samples", "config1_small_lp.cc"))], capture_output=True, text=True, timeout=300)
    assert run.returncode == 0, (run.returncode, run.stdout, run.stderr)
    lines = run.stdout.splitlines()
    assert lines[0].startswith("TERMINATION_REASON_OPTIMAL after ")
    assert abs(float(lines[1].split()[2]) + 34.0) < 1e-4
    primal = [float(line.split()[1]) for line in lines if line.startswith("x")]
    assert max(abs(a - b) for a, b in zip(primal, [-1, 8, 1, 2.5])) < 1e-4  # test_util.h:45-48

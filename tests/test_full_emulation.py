"""The whole library on CPU - host driver AND kernels - as a test of the host driver's logic (`-m "not gpu"`).

tests/emu_lib/build.py compiles every source of or-tools_b200/csrc with g++ as host C++ against a stand-in for the
CUDA runtime (tests/emu_lib/include/cuda_runtime.h: one device, synchronous streams, device memory = host memory);
kernel launches become calls that run each CTA as 256 user-level contexts (tests/kernel_emu/cuda_emu.h, fiber mode).
The result, tests/_build/libpdlp_b200_emu.so, has the C ABI of the product and is loaded INSTEAD of it by a child
pytest process (PDLP_B200_LIBRARY) that runs a selection of the `-m gpu` tests: small problems only, an emulated launch
costs milliseconds.

This is test infrastructure, not a CPU path of the product: the package never builds, ships or picks this library, and
without PDLP_B200_LIBRARY pointing at it a solve on a machine without a GPU fails loudly (tests/test_cabi.py).  It
checks what the kernel-level emulation cannot: launch order, buffer sizes, the device-resident loop's state hand-over,
the degenerate shapes, the presolve hand-off - everything except timing, asynchrony and the hardware memory model."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "emu_lib"))


@pytest.fixture(scope="module")
def emulated_library():
    import build  # tests/emu_lib/build.py

    return build.build()


def run_device_tests(library, selection, timeout):
    env = dict(os.environ, PDLP_B200_LIBRARY=library)
    cmd = [sys.executable, "-m", "pytest", "-q", "-m", "gpu", "-p", "no:cacheprovider", "--timeout", "120"] + selection
    run = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=timeout)
    tail = "\n".join(run.stdout.splitlines()[-25:])
    assert run.returncode == 0, tail + "\n" + run.stderr[-2000:]
    return tail


def test_the_emulated_library_is_not_the_product(emulated_library):
    import pdlp_b200

    assert os.path.realpath(emulated_library) != os.path.realpath(pdlp_b200.LIBRARY_PATH)
    assert os.path.commonpath([emulated_library, os.path.join(ROOT, "tests")]) == os.path.join(ROOT, "tests")
    assert not os.path.exists(os.path.join(ROOT, "or-tools_b200", "lib", os.path.basename(emulated_library)))


def test_degenerate_shapes_and_presolve_rules_through_the_host_driver(emulated_library):
    """Problems without rows / without columns, the with-presolve small-value checks, a presolver that solves everything,
    the presolve goldens: device tests that were written after the last run on a B200."""
    tail = run_device_tests(emulated_library, ["tests/test_zy_late_additions.py", "tests/test_presolve.py",
                                               "-k", "not concurrent and not plain_c and not reports_on_the_original"], 600)
    assert " passed" in tail and "failed" not in tail


def test_reference_behaviour_and_unit_tests_through_the_host_driver(emulated_library):
    """The reference's behavioural tests (restart bookkeeping, check cadence, callbacks, warm starts, interrupts, every
    Detects* case, feasibility polishing, all three step rules) and its unit goldens (statistics, rescaling, KKT passes,
    reduced costs, localized bounds) in their [device] variants: about 100 solves of the reference's small fixtures."""
    tail = run_device_tests(emulated_library, ["tests/test_reference_pdhg_behaviour.py", "tests/test_reference_units.py", "tests/test_logging.py"], 1200)
    assert " passed" in tail and "failed" not in tail


def test_python_surface_solves_through_the_host_driver(emulated_library):
    """The reference's Python tests (pdlp_test.py) and the config-1 sample on the emulated device: tight-tolerance solves of
    TinyLp / TestLp (hundreds of iterations of the device-resident loop), a warm start at the optimum, an iteration limit."""
    tail = run_device_tests(emulated_library, ["tests/test_python_surface.py"], 900)
    assert " passed" in tail and "failed" not in tail


def test_c_and_cpp_callers_in_device_mode(emulated_library):
    """tests/c/caller.c, tests/cpp/caller.cc and samples/config1_small_lp.cc linked against the emulated library and run
    in the mode the GPU box runs them in (a TestLp solve through pdlp_b200_solve / PrimalDualHybridGradient)."""
    out_dir = os.path.join(ROOT, "tests", "_build", "emu_callers")
    os.makedirs(out_dir, exist_ok=True)
    libdir, libname = os.path.dirname(emulated_library), os.path.basename(emulated_library)[3:-3]
    link = ["-L" + libdir, "-l" + libname, "-Wl,-rpath," + libdir]
    include = "-I" + os.path.join(ROOT, "include")
    builds = [
        ("c_caller", ["gcc", "-std=c99", include, os.path.join(ROOT, "tests", "c", "caller.c")] + link + ["-lm"], ["gpu"], "gpu ok"),
        ("cpp_caller", ["g++", "-std=c++17", include, os.path.join(ROOT, "tests", "cpp", "caller.cc")] + link, ["gpu"], "gpu ok: OPTIMAL"),
        ("config1_small_lp", ["g++", "-std=c++17", include, os.path.join(ROOT, "samples", "config1_small_lp.cc")] + link, [], "TERMINATION_REASON_OPTIMAL after "),
    ]
    for name, compile_cmd, args, expected in builds:
        exe = os.path.join(out_dir, name)
        subprocess.run(compile_cmd + ["-o", exe], check=True)
        run = subprocess.run([exe] + args, capture_output=True, text=True, timeout=300)
        assert run.returncode == 0 and expected in run.stdout, (name, run.returncode, run.stdout, run.stderr)

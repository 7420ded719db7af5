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
the degenerate shapes, the presolve hand-off - everything except timing, asynchrony and the hardware memory model.

Also here: column-sharded solves whose ranks are host threads of one process (peer memory = plain memory, NCCL = a thread
rendezvous, tests/emu_lib/fake_nccl.cc) at 2, 4 and 8 ranks; the per-iteration interrupt; a rank that cannot build its
block; every device allocation of a solve failing in turn.  Sanitizer variants of the same library (AddressSanitizer,
ThreadSanitizer with one host thread per CUDA thread) are run by scripts/emulated_asan.sh / emulated_tsan.sh, not here:
they take the better part of an hour (logs under profiles/)."""
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


# Groups of `-m gpu` tests run on the emulated library, each in its own child pytest, all at the same time (an emulated
# solve is single-threaded).  Left out: problems of thousands of rows (minutes per solve at ~1 ms per emulated launch) and
# the test that times a flag raised 0.25 s into a solve against the wall clock.
GROUPS = {
    "degenerate_shapes_and_presolve": ["tests/test_zy_late_additions.py", "tests/test_presolve.py", "-k", "not concurrent and not plain_c and not reports_on_the_original"],
    "reference_behaviour": ["tests/test_reference_pdhg_behaviour.py", "tests/test_logging.py"],
    "reference_units": ["tests/test_reference_units.py", "tests/test_gpu_trial_steps.py", "tests/test_gpu_forced_termination.py"],
    "python_surface_and_boundary": ["tests/test_python_surface.py", "tests/test_qp_io.py", "tests/test_mp_model.py", "tests/test_cabi.py", "tests/test_solve_cli.py"],
    "device_against_oracle": ["tests/test_gpu_solve.py", "-k", "not test_random_lp_parity"],
    # Two of the reference's sixteen parameter combinations over its 12 LP and 4 QP tests (all 320 [device] cases pass on
    # the emulated library - 12 minutes in one process - with `pytest tests/test_reference_pdhg.py -m gpu` under the same
    # PDLP_B200_LIBRARY).
    "reference_parameter_combinations": ["tests/test_reference_pdhg.py", "-k", "device-Scaling_1T_IterLimit_Adaptive or device-NoScaling_4T_NoIterLimit_Adaptive"],
}


@pytest.fixture(scope="module")
def device_runs(emulated_library):
    env = dict(os.environ, PDLP_B200_LIBRARY=emulated_library)
    children = {}
    for name, selection in GROUPS.items():
        cmd = [sys.executable, "-m", "pytest", "-q", "-m", "gpu", "-p", "no:cacheprovider", "--timeout", "300"] + selection
        children[name] = subprocess.Popen(cmd, cwd=ROOT, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    results = {}
    for name, child in children.items():
        try:
            out, _ = child.communicate(timeout=1500)
        except subprocess.TimeoutExpired:
            child.kill()
            out, _ = child.communicate()
            out += "\n(killed after 1500 s)"
        results[name] = (child.returncode, "\n".join(out.splitlines()[-30:]))
    return results


def test_the_emulated_library_is_not_the_product(emulated_library):
    import pdlp_b200

    assert os.path.realpath(emulated_library) != os.path.realpath(pdlp_b200.LIBRARY_PATH)
    assert os.path.commonpath([emulated_library, os.path.join(ROOT, "tests")]) == os.path.join(ROOT, "tests")
    assert not os.path.exists(os.path.join(ROOT, "or-tools_b200", "lib", os.path.basename(emulated_library)))


@pytest.mark.parametrize("group", sorted(GROUPS))
def test_device_tests_through_the_emulated_host_driver(device_runs, group):
    """degenerate_shapes_and_presolve: problems without rows / columns, the with-presolve small-value checks, a presolver
    that solves everything - written after the last run on a B200.  reference_behaviour / reference_units: the reference's
    behavioural tests (restarts, cadence, callbacks, warm starts, interrupts, Detects*, polishing, three step rules) and unit
    goldens in their [device] variants.  python_surface_and_boundary: pdlp_test.py, the sample, file readers, MPModelProto
    requests, the C ABI, the command line.  device_against_oracle: iteration traces and objectives against the oracle.
    reference_parameter_combinations: the reference's LP / QP solve tests under two of the adaptive-rule combinations (the Malitsky-Pock ones take minutes here; that rule runs in device_against_oracle)."""
    code, tail = device_runs[group]
    assert code == 0 and " passed" in tail and " failed" not in tail, tail


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


# ---- the column-sharded solve with ranks as host threads (peer-memory path at 2, 4 and 8 ranks; NCCL path at 2) ----

SHARDED = [(2, "peer", "random"), (4, "peer", "random"), (8, "peer", "random"), (2, "nccl", "random"), (2, "peer", "flow"), (8, "peer", "flow"), (4, "peer", "qp_mp")]


@pytest.fixture(scope="module")
def sharded_runs(emulated_library):
    import json

    import build  # tests/emu_lib/build.py

    env = dict(os.environ, PDLP_B200_LIBRARY=emulated_library, PDLP_B200_NCCL_LIBRARY=build.build_fake_nccl())
    script = os.path.join(ROOT, "tests", "emu_lib", "sharded_threads.py")
    children = {cfg: subprocess.Popen([sys.executable, script, str(cfg[0]), cfg[1], cfg[2]], cwd=ROOT, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
                for cfg in SHARDED}
    results = {}
    for cfg, child in children.items():
        try:
            out, err = child.communicate(timeout=900)
        except subprocess.TimeoutExpired:  # a rank waiting for a flag that never comes: the hang a real run would show
            child.kill()
            out, err = child.communicate()
            err += "\n(killed after 900 s)"
        lines = [line for line in out.splitlines() if line.startswith("{")]
        results[cfg] = (child.returncode, json.loads(lines[-1]) if lines else None, err[-3000:])
    return results


@pytest.mark.parametrize("world, mode, problem", SHARDED)
def test_sharded_solve_with_ranks_as_host_threads(sharded_runs, world, mode, problem):
    """The shipped exchange paths through the whole host driver: communicator and peer-region set-up (IPC handles through an
    all-reduce, the agreement on failure), the fused peer pass with its CTA split and in-kernel signal, the finishing kernel,
    the scalar exchange inside the step-decision kernel, host events with their collectives, the gather of the result.  Peer
    memory is plain memory here and NCCL a thread rendezvous (tests/emu_lib/fake_nccl.cc), so this says nothing about NVLink
    ordering - the 2-GPU tests on the B200 box do - but every rank must finish, agree, and match the one-rank solve.
    "random": nearly every row shared; "flow": config 3 in miniature, private conservation rows and shared capacity rows, the
    layout in which the fused pass splits its grid and signals from inside the kernel; "qp_mp": a diagonal QP under the
    Malitsky-Pock rule (its own decision kernel and two-sided averaging)."""
    code, got, err = sharded_runs[(world, mode, problem)]
    assert code == 0 and got is not None, err
    optimal = int(__import__("pdlp_b200").TerminationReason.TERMINATION_REASON_OPTIMAL)
    assert got["single_reason"] == optimal and got["reasons"] == [optimal] * world
    assert len(set(got["iterations"])) == 1  # the ranks stop at the same iteration
    assert 0.5 <= got["iterations"][0] / got["single_iterations"] <= 2.0, got
    assert abs(got["objective_sharded"] - got["objective_single"]) <= 10 * got["eps"] * abs(got["objective_single"]), got  # both eps-optimal
    if problem == "flow":
        assert min(got["private_rows"]) > 0 and 0 < got["shared_rows"] < got["rows"]
    assert got["x_distance"] < 1e-2 and got["y_distance"] < 1e-3 and got["duals_identical_on_shared_rows"]
    assert len(set(got["collectives"])) == 1 and got["collectives"][0] > got["iterations"][0]  # at least one exchange per trial step
    if mode == "peer":
        assert len(set(got["peer_exchanges"])) == 1 and got["peer_exchanges"][0] > got["iterations"][0], got
    else:
        assert got["peer_exchanges"] == [0] * world


def test_peer_and_nccl_exchange_give_the_same_bits(sharded_runs):
    """Both paths add the ranks' row sums in rank order: the solves must agree to the last bit."""
    peer, nccl = sharded_runs[(2, "peer", "random")][1], sharded_runs[(2, "nccl", "random")][1]
    assert peer is not None and nccl is not None
    assert peer["iterations"] == nccl["iterations"] and peer["objective_sharded"] == nccl["objective_sharded"]
    assert peer["x_distance"] == nccl["x_distance"] and peer["y_distance"] == nccl["y_distance"]


def test_interrupt_raised_mid_solve_is_seen_by_the_decision_kernel(emulated_library):
    """tests/emu_lib/interrupt_mid_solve.py: host events 4096 iterations apart, the flag raised 2 s into the solve from another
    thread.  The loop must stop between host events, and the counter must say that the decision kernel saw the flag."""
    import json

    env = dict(os.environ, PDLP_B200_LIBRARY=emulated_library)
    run = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "emu_lib", "interrupt_mid_solve.py")], cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    assert run.returncode == 0, run.stderr[-2000:]
    got = json.loads([line for line in run.stdout.splitlines() if line.startswith("{")][-1])
    assert got["reason"] == int(__import__("pdlp_b200").TerminationReason.TERMINATION_REASON_INTERRUPTED_BY_USER)
    assert got["seen_on_device"] == 1, got
    assert 0 < got["iterations"] < 4096, got  # before the first host event after iteration 0's
    assert got["latency"] < 5.0, got  # at most two 32-trial chunks are queued behind the running one, also here


def test_a_rank_that_cannot_build_its_block_fails_every_rank(emulated_library):
    """Rank 3 of 4 holds a block that is not a valid CSC matrix.  It must report that, and the other ranks must come back
    with an error of their own instead of waiting in the next collective for a rank that has given up."""
    import json

    import build  # tests/emu_lib/build.py

    env = dict(os.environ, PDLP_B200_LIBRARY=emulated_library, PDLP_B200_NCCL_LIBRARY=build.build_fake_nccl())
    run = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "emu_lib", "sharded_threads.py"), "4", "peer", "bad_block"], cwd=ROOT, env=env,
                         capture_output=True, text=True, timeout=300)  # a hang ends here
    assert run.returncode == 0, run.stderr[-2000:]
    got = json.loads([line for line in run.stdout.splitlines() if line.startswith("{")][-1])
    assert got["returned_a_result"] == [False] * 4 and len(got["errors"]) == 4
    by_rank = {e.split(":")[0]: e for e in got["errors"]}
    assert "not a valid compressed CSC matrix" in by_rank["rank 3"]
    assert all("another rank could not build its block" in by_rank[f"rank {r}"] for r in range(3))


def test_every_device_allocation_of_a_solve_may_fail(emulated_library):
    """tests/emu_lib/allocation_failures.py: each of the solve's device allocations reports out-of-memory once, in turn.
    Every attempt must return the ABI's CUDA-error status (a RuntimeError here), and the next solve must be untouched."""
    import json

    env = dict(os.environ, PDLP_B200_LIBRARY=emulated_library)
    run = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "emu_lib", "allocation_failures.py")], cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    assert run.returncode == 0, run.stdout[-1000:] + run.stderr[-2000:]
    got = json.loads([line for line in run.stdout.splitlines() if line.startswith("{")][-1])
    assert got["allocations_per_solve"] >= 20 and got["failed_cleanly"] == got["allocations_per_solve"]
    assert all(m.startswith("pdlp_b200 status 2: CUDA error") for m in got["messages"]) and got["solve_after_failures_identical"]

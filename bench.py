#!/usr/bin/env python
"""Benchmark of the PDLP hot path (BASELINE.json metric): PDHG iterations/sec, time to
1e-4 relative KKT, achieved HBM GB/s.

    python bench.py --gpus N --steps K --warmup W [--workload c2|c3|c4|c5] [--impl reference]

A "step" is one PDHG trial step (primal step, K x fused with the dual step, K^T y fused
with the step-size test, decision, averaging) on the resident problem.
  value     trial steps / second, problem and iterate resident in HBM, CUDA-event timed
  e2e       iterations / second of a full solve to eps through the public API with HOST buffers
            (upload, transpose, rescaling, termination checks, restarts, download included)
  loop_iterations_per_sec / loop_roofline_frac
            the same solve's iteration_count / (solve time - preprocessing) and the
            algorithmic bytes of its trial steps over that time against the HBM peak
  roofline  algorithmic bytes per step / measured step time vs the
            measured HBM copy bandwidth in MEASURED_PEAKS.json, plus a per-kernel table
  parity    the CPU oracle's (tests/oracle_binding.py) convergence information at the device's
            solution: relative KKT errors <= eps, objective against the planted optimum where the
            generator plants one, iteration count against the N = 1 solve of the same LP
  cpu_baseline  the CPU oracle's step loop (a restatement of the reference's threaded Sharder path)
            over a bounded number of steady-state iterations: the same unit as `value`
`--impl reference` times that CPU path alone, without loading the CUDA library.  Multi-GPU (N > 1):
the SAME problem column-sharded over the N ranks (strong scaling): one process per GPU,
each owning a column block of K.  `--replicas` runs N independent copies instead (no collective).
"""
import argparse
import importlib
import json
import math
import os
import signal
import subprocess
import sys
import tempfile
import threading
import time
import types

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
    os.environ["NCCL_DEBUG"] = "NONE"  # NCCL's version banner (level VERSION, e.g. from /etc/nccl.conf) goes to stdout, which carries the one JSON line

import numpy as np  # noqa: E402

WORKLOADS = {
    "c2": "synthetic random sparse LP 1M rows x 2M cols, 1e7 nnz, box bounds, eps=1e-4",
    "c3": "synthetic multi-commodity network-flow LP, 64 commodities x 65536 nodes x 520000 arcs, ~1e8 nnz, block-angular order (8 blocks of 8 commodities, arc-major inside a block), eps=1e-4",
    "c3a": "the c3 LP in arc-major order (a contiguous column block is a set of arcs and touches nearly every row), eps=1e-4",
    "c3k": "the c3 LP in commodity-major order (textbook block-angular, strided capacity rows), eps=1e-4",
    "c4": "synthetic diagonal-objective convex QP 5M rows x 10M cols, 5e7 nnz, eps=1e-6",
    "c5": "synthetic transportation LP (dense bipartite), 1.25e8 nnz at scale 1 (scale 8 = the 1e9-nnz LP of config 5), eps=1e-4",
}
# Steady-state iterations of the CPU step loop in the cpu_baseline leg / the reference arm (plus 2 untimed).
CPU_STEPS = {"c2": 64, "c3": 16, "c3a": 16, "c3k": 16, "c4": 24, "c5": 12}


def synthetic_module(native):
    """or-tools_b200/synthetic.py.  native=False loads it (and the library-free QuadraticProgram of
    _qp.py) under a stand-in package name WITHOUT executing the package's __init__, which would load
    libpdlp_b200.so: the reference arm must not map the product library."""
    if native:
        from pdlp_b200 import synthetic
        return synthetic
    name = "pdlp_b200_hostonly"
    if name not in sys.modules:
        pkg = types.ModuleType(name)
        pkg.__path__ = [os.path.join(ROOT, "or-tools_b200")]
        sys.modules[name] = pkg
    return importlib.import_module(name + ".synthetic")


def make_problem(name, scale, seed=0, native=True):
    synthetic = synthetic_module(native)
    if name == "c2":
        return synthetic.random_sparse_lp(int(1_000_000 * scale), int(2_000_000 * scale), 5, seed)
    if name in ("c3", "c3a", "c3k"):
        side = max(8, int(round(256 * scale ** 0.5)))
        return synthetic.multicommodity_flow_lp(64, side, int(520000 * scale), 16, seed, order={"c3": "block", "c3a": "arc", "c3k": "commodity"}[name])
    if name == "c4":
        return synthetic.random_sparse_lp(int(5_000_000 * scale), int(10_000_000 * scale), 5, seed, qp_diagonal=True)
    if name == "c5":
        s = c5_side(scale)
        return synthetic.transportation_lp(s, s, seed)
    raise ValueError(name)


def c5_side(scale):
    return max(8, int(round(7906 * scale ** 0.5)))  # 7906^2 * 2 = 1.25e8 nnz at scale 1; scale 8 is the 1e9-nnz LP of config 5


def problem_dims(name, scale):
    """(n, m, nnz) of the whole workload without generating it (config of the reference arm at any N)."""
    if name in ("c2", "c4"):
        base_m, base_n = (1_000_000, 2_000_000) if name == "c2" else (5_000_000, 10_000_000)
        m, n = int(base_m * scale), int(base_n * scale)
        return n, m, n * min(5, m)
    if name == "c5":
        s = c5_side(scale)
        return s * s, 2 * s, 2 * s * s
    return None


def make_block(name, scale, rank, world, seed=0):
    """This rank's column block of the workload in the sharded solver's local constraint order:
    (block qp, column starts, number of shared constraints, global constraint count, whole qp or
    None, RowSharing or None).  c5 is generated block by block (whole sources, so the 1e9-nnz problem
    never exists on one host); the others are generated whole on every rank (same seed) and cut by
    the mass rule."""
    import pdlp_b200
    from pdlp_b200 import synthetic
    if name == "c5":
        s = c5_side(scale)
        cuts = [(s * r) // world for r in range(world + 1)]
        block = synthetic.transportation_lp(s, s, seed, cuts[rank], cuts[rank + 1], local_rows=True)
        return block, np.array([c * s for c in cuts], dtype=np.int64), s, 2 * s, None, None
    qp = make_problem(name, scale, seed)
    K = qp.constraint_matrix
    starts = pdlp_b200.column_partition(K.indptr, world)
    sharing = pdlp_b200.RowSharing.from_matrix(K, starts)
    block = pdlp_b200.take_columns(qp, int(starts[rank]), int(starts[rank + 1]), sharing.local_rows(rank))
    return block, starts, int(sharing.shared.size), K.shape[0], qp, sharing


def workload_eps(name):
    return 1e-6 if name == "c4" else 1e-4


def solve_params(name, threads=1, iteration_limit=None):
    import pdlp_b200
    p = pdlp_b200.PrimalDualHybridGradientParams()
    eps = workload_eps(name)
    p.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = eps
    p.termination_criteria.simple_optimality_criteria.eps_optimal_relative = eps
    p.num_threads = threads
    # Guard rail of the end-to-end leg only (a solve that reaches eps never sees it; the reason it
    # stopped for is reported in e2e.termination_reason): the reference's default is unlimited.  No
    # time limit: a finite one makes the host poll the clock every ~20 ms instead of every 64 iterations.
    p.termination_criteria.iteration_limit = 100000 if iteration_limit is None else iteration_limit
    return p


def oracle_params_text(name, threads, iteration_limit=None):
    """The same parameters as solve_params() in the oracle's 'field=value' lines (tests/oracle_binding.py
    flatten_params), written by hand so that the reference arm needs no class of the product package."""
    eps = repr(float(workload_eps(name)))
    lines = ["termination_criteria=", "termination_criteria.simple_optimality_criteria=",
             "termination_criteria.simple_optimality_criteria.eps_optimal_absolute=" + eps,
             "termination_criteria.simple_optimality_criteria.eps_optimal_relative=" + eps,
             "termination_criteria.iteration_limit=" + str(100000 if iteration_limit is None else iteration_limit),
             "num_threads=" + str(threads)]
    return "\n".join(lines)


def bytes_per_step(n, m, nnz, is_qp, sharded=False):
    """B_step = 24 nnz + 4 (n + m) + 8 (15 n + 8 m)  [+ 8 n for a QP].
    Sharded: the same for one rank's block (n, nnz and m of the block: its shared + private
    constraints) plus the partial row sums written before and read after the exchange (16 m);
    the exchange's own traffic is not counted."""
    return 24 * nnz + 4 * (n + m) + 8 * (15 * n + 8 * m) + (8 * n if is_qp else 0) + (16 * m if sharded else 0)


def kernel_bytes(n, m, nnz, is_qp):
    """Algorithmic bytes of each kernel of one accepted trial step.  The weighted
    average of SURVEY 8(d) (3n + 3m doubles) is folded into the next step's K1 (x part) and K2 (y part),
    so its algorithmic bytes are counted there; the "average" slot is the empty interval left in the
    per-kernel profile (the batch-end flush is outside it)."""
    return {
        "primal_step": 8 * (7 * n) + (8 * n if is_qp else 0) + 8 * 3 * n,
        "dual_step_Kx": 12 * nnz + 4 * m + 8 * (n + 4 * m) + 8 * 3 * m,
        "kty_step_KTy": 12 * nnz + 4 * n + 8 * (4 * n + m),
        "decide": 0,
        "average": 0,
    }


def make_config(name, scale, n, m, nnz, world, sharded):
    """One dict for both arms (`--impl reference` prints the same keys and values)."""
    return {"workload": name + ": " + WORKLOADS[name], "scale": scale, "seed": 0,
            "l2_policy": "inputs larger than L2 (matrix + vectors >> 126 MB); no flush between steps",
            "n": int(n), "m": int(m), "nnz": int(nnz),
            "parallelism": "1 GPU" if world == 1 else (f"column-sharded over {world} GPUs; the communicator and its peer-memory mappings are created by the first "
                                                         "sharded call of the process (the untimed resident problem) and reused by the end-to-end solve" if sharded
                                                         else f"{world} independent replicas")}


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""

    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device = device
        self.samples = []
        self.stop = threading.Event()
        self.thread = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits", "-i", str(self.device)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([t.strip() for t in out.splitlines()[0].split(",")])
            except Exception:  # noqa: BLE001
                pass
            self.stop.wait(0.1)

    def __enter__(self):
        self.thread.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.thread.join(timeout=6)

    def summary(self):
        sm = [float(s[1]) for s in self.samples if len(s) >= 8 and s[1].replace(".", "").isdigit()]
        mx = [float(s[2]) for s in self.samples if len(s) >= 8 and s[2].replace(".", "").isdigit()]
        reasons = set()
        for s in self.samples:
            if len(s) >= 8:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), s[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def recorded_traffic(workload, kernel):
    """DRAM bytes per launch of `kernel` from the committed ncu capture (profiles/traffic.json), or None."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(path) as f:
            return json.load(f).get(workload, {}).get(kernel)
    except (OSError, ValueError):
        return None


def cpu_step_leg(qp, name, threads, steps, warmup=2):
    """Trial steps / second of the CPU oracle's step loop in its steady state (threaded Sharder
    structure, num_shards = 4 x threads; `warmup` untimed + `steps` timed iterations of TakeAdaptiveStep,
    rejected trials counted like the device's; no termination checks, no restarts): the unit of `value`."""
    import oracle_binding as ob
    r = ob.time_steps(qp, oracle_params_text(name, threads), warmup, steps)
    trials = r["accepted_steps"] + r["rejected_steps"]
    return {"value": trials / max(r["seconds"], 1e-9), "trial_steps": trials, "accepted_steps": r["accepted_steps"], "seconds": r["seconds"],
            "preprocessing_seconds": r["preprocessing_seconds"]}


def cpu_loop_leg(qp, name, threads, base, extra):
    """Iterations / second of the oracle's whole solve loop (termination checks every 64 iterations,
    restart evaluations, everything primal_dual_hybrid_gradient.cc:2689-2774 does) by a difference of
    two solves with iteration limits `base` and `base + extra`: the first iterations are dominated by
    the artificial restarts at 1, 2, 3, 5, 9, 17, 33 (:1915-1925) and the check at iteration 0, which a
    solve of hundreds of iterations amortises."""
    import oracle_binding as ob
    t = []
    for limit in (base, base + extra):
        log = ob.solve(qp, oracle_params_text(name, threads, limit)).solve_log_dict
        if log["iteration_count"] != limit:
            return None  # converged before the limit: nothing to difference
        t.append(log["solve_time_sec"] - log["preprocessing_time_sec"])
    return {"value": extra / max(t[1] - t[0], 1e-9), "base_iterations": base, "extra_iterations": extra, "loop_seconds": t}


def relative_kkt(stats, eps, bound_norm, objective_norm):
    """The three optimality tests of termination.cc:26-100 (L2 norm) as ratios error / allowance."""
    gap = abs(stats["primal_objective"] - stats["dual_objective"])
    return {"primal_residual": stats["l2_primal_residual"] / (eps + eps * bound_norm),
            "dual_residual": stats["l2_dual_residual"] / (eps + eps * objective_norm),
            "objective_gap": gap / (eps + eps * (abs(stats["primal_objective"]) + abs(stats["dual_objective"])))}


def compare_with_n1(out, name, scale, world, iterations, objective, eps):
    """N = 1: remembers (iterations, objective) of this workload in the temp directory.  N > 1: adds out["vs_n1"] when this
    box has run the N = 1 solve of the same LP; returns False when the two objectives differ by more than 4 eps (two
    eps-optimal points of one LP)."""
    ref_path = os.path.join(tempfile.gettempdir(), f"pdlp_b200_bench_n1_{name}_s{scale:g}.json")
    if world == 1:
        try:
            with open(ref_path, "w") as f:
                json.dump({"iterations": int(iterations), "primal_objective": objective}, f)
        except OSError:
            pass
    elif os.path.exists(ref_path):
        with open(ref_path) as f:
            ref = json.load(f)
        out["vs_n1"] = {"iterations_n1": ref["iterations"], "iteration_ratio": iterations / max(1, ref["iterations"]),
                        "objective_rel_diff": abs(objective - ref["primal_objective"]) / (1.0 + abs(objective))}
        return out["vs_n1"]["objective_rel_diff"] <= 4 * eps
    return True


def parity_check(qp, name, scale, x, y, device_objective, iterations, world):
    """The oracle's view of the device's solution (rank 0, whole problem): ComputeConvergenceInformation
    (iteration_stats.cc:349-419) at (x, y) on the ORIGINAL problem; every relative KKT error must be
    within the requested eps (1e-6 relative slack for the different summation order), the device's own
    objective must agree with the oracle's at that point, and where the generator planted an optimum
    the objective must match it to the accuracy eps implies.  N > 1 also reports the iteration count
    against the N = 1 solve of the same LP when this box has run it (a file in the temp directory)."""
    import oracle_binding as ob
    eps = workload_eps(name)
    t0 = time.time()
    prob = ob.Problem(qp, num_threads=min(16, os.cpu_count() or 1), num_shards=4 * min(16, os.cpu_count() or 1))
    st = prob.stats()
    stats = prob.convergence_info(oracle_params_text(name, 1), np.ones(len(x)), np.ones(len(y)), x, y)
    ratios = relative_kkt(stats, eps, st["combined_bounds_l2_norm"], st["objective_vector_l2_norm"])
    out = {"checker": "CPU oracle ComputeConvergenceInformation at the device's (x, y) on the unscaled problem", "eps": eps,
           "relative_kkt_over_eps": ratios, "oracle_primal_objective": stats["primal_objective"], "oracle_dual_objective": stats["dual_objective"],
           "device_primal_objective": device_objective, "iterations": int(iterations)}
    ok = all(v <= 1.0 + 1e-6 for v in ratios.values())
    scale = 1.0 + abs(stats["primal_objective"])
    out["objective_vs_oracle_rel"] = abs(device_objective - stats["primal_objective"]) / scale
    ok = ok and out["objective_vs_oracle_rel"] <= 1e-9
    planted = getattr(qp, "planted_primal", None)
    if planted is not None:
        opt = float(qp.objective_vector @ planted)
        if qp.objective_matrix is not None:
            opt += 0.5 * float(planted @ (qp.objective_matrix * planted))
        # |objective - optimum| <= gap allowance + |y*| x primal residual allowance, generously: 10 eps (1 + |p| + |d|)
        out["planted_objective"] = opt
        out["objective_vs_planted_rel"] = abs(stats["primal_objective"] - opt) / (1.0 + abs(opt))
        out["objective_vs_planted_tolerance"] = 20 * eps
        ok = ok and out["objective_vs_planted_rel"] <= 20 * eps
    ok = compare_with_n1(out, name, scale, world, iterations, stats["primal_objective"], eps) and ok
    out["ok"] = bool(ok)
    out["seconds"] = time.time() - t0
    return out


def transportation_parity(qp, x, y, local_rows, device_objective, iterations, reduce_sum):
    """ComputeConvergenceInformation (iteration_stats.cc:349-419) restated in numpy for the structure of
    the transportation LP (synthetic.transportation_lp): K x is a row sum and a column sum of x seen as a
    (sources, sinks) table, K^T y is y_supply[s] + y_demand[d].  For the c5 lines whose matrix never
    exists on one host (sharded) or is too large for a second copy in the oracle (1e9 nnz): every rank
    passes its own block (whole sources; `local_rows` layout: sink rows first, then its supply rows) and
    `reduce_sum` adds small vectors over the ranks.  tests/test_bench_helpers.py pins it to the oracle."""
    cub = np.asarray(qp.constraint_upper_bounds)
    m, n = len(cub), len(x)
    # Rows: local_rows -> [D sink rows, the block's supply rows]; whole problem -> [S supply rows, D sink rows].  The sink rows are the equalities.
    finite = np.isfinite(np.asarray(qp.constraint_lower_bounds))
    if local_rows:
        D = int(finite.sum())  # the sink rows are the equalities
        nblock = m - D
        demand, supply = cub[:D], cub[D:]
        y_d, y_s = y[:D], y[D:]
    else:
        nblock = int((~finite).sum())
        D = m - nblock
        supply, demand = cub[:nblock], cub[nblock:]
        y_s, y_d = y[:nblock], y[nblock:]
    assert D * nblock == n, (D, nblock, n, m)
    X = np.asarray(x).reshape(nblock, D)
    c = np.asarray(qp.objective_vector).reshape(nblock, D)
    sink_activity = reduce_sum(X.sum(axis=0))
    supply_violation = np.maximum(X.sum(axis=1) - supply, 0.0)
    # Reduced costs c - K^T y; x >= 0 only: a negative reduced cost is dual residual, a positive one pairs with the bound 0.
    dual_res2 = 0.0
    chunk = max(1, (1 << 24) // max(1, D))
    for a in range(0, nblock, chunk):
        r = c[a:a + chunk] - y_d[None, :] - y_s[a:a + chunk, None]
        dual_res2 += float(np.square(np.minimum(r, 0.0)).sum())
    # Supply rows have no lower bound: a positive dual there makes the dual objective -inf (iteration_stats.cc:294-316); PDHG
    # iterates are projected onto y_s <= 0, so it does not occur at a solver's point.
    sums = reduce_sum(np.array([float(np.square(supply_violation).sum()), dual_res2, float((c * X).sum()), float((supply * np.minimum(y_s, 0.0)).sum()),
                                float(np.square(supply).sum()), float(np.square(c).sum())]))
    primal_res = math.sqrt(sums[0] + float(np.square(sink_activity - demand).sum()))
    dual_res = math.sqrt(sums[1])
    primal_obj = sums[2]
    dual_obj = sums[3] + float(demand @ y_d)
    if reduce_sum(np.array([float((y_s > 0.0).sum())]))[0] > 0:
        dual_obj = -math.inf
    eps = 1e-4
    bound_norm = math.sqrt(sums[4] + float(np.square(demand).sum()))
    ratios = relative_kkt({"primal_objective": primal_obj, "dual_objective": dual_obj, "l2_primal_residual": primal_res, "l2_dual_residual": dual_res}, eps,
                          bound_norm, math.sqrt(sums[5]))
    out = {"checker": "numpy restatement of ComputeConvergenceInformation on the transportation structure (row / column sums of the (sources, sinks) table), "
                      "each rank its own block, pinned to the CPU oracle by tests/test_bench_helpers.py", "eps": eps, "relative_kkt_over_eps": ratios,
           "checker_primal_objective": primal_obj, "checker_dual_objective": dual_obj, "device_primal_objective": device_objective, "iterations": int(iterations)}
    out["objective_vs_checker_rel"] = abs(device_objective - primal_obj) / (1.0 + abs(primal_obj))
    out["ok"] = bool(all(v <= 1.0 + 1e-6 for v in ratios.values()) and out["objective_vs_checker_rel"] <= 1e-9)
    return out


def reference_arm(args, name, world):
    """`--impl reference`: the CPU oracle (the restatement of the reference's own CPU path; the reference
    itself cannot be built here) on the host cores.  Imports nothing of the product package."""
    threads = os.cpu_count() or 1
    qp = make_problem(name, args.scale, native=False)
    K = qp.constraint_matrix
    steps = args.cpu_steps or CPU_STEPS[name]
    for _ in range(min(args.warmup, 1)):
        cpu_step_leg(qp, name, threads, 1, warmup=0)
    res = cpu_step_leg(qp, name, threads, steps)
    loop = None if args.no_cpu_loop else cpu_loop_leg(qp, name, threads, 64, 64)
    sample = (f"{res['trial_steps']} steady-state trial steps ({res['accepted_steps']} accepted) of the oracle's step loop after 2 untimed ones on the full workload "
              f"(n={K.shape[1]}, m={K.shape[0]}, nnz={K.nnz}); no termination checks or restarts in the sample; {res['seconds']:.2f} s, preprocessing {res['preprocessing_seconds']:.2f} s excluded")
    sharded = world > 1 and not args.replicas
    line = {
        "impl": "reference", "metric": "pdhg_iterations_per_sec", "value": res["value"], "unit": "iterations/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 / res["value"], "higher_is_better": True, "scaling": "strong" if sharded or world == 1 else "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": make_config(name, args.scale, K.shape[1], K.shape[0], K.nnz, world, sharded),
        "cpu_baseline": {"value": res["value"], "unit": "iterations/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": res["value"], "unit": "iterations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "loop_iterations_per_sec": None if loop is None else loop["value"], "loop_sample": loop,
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=os.environ.get("PDLP_BENCH_WORKLOAD", "c3"), choices=sorted(WORKLOADS))
    ap.add_argument("--scale", type=float, default=float(os.environ.get("PDLP_BENCH_SCALE", "1.0")), help="shrinks or grows the workload (c5 at 8 = the 1e9-nnz LP)")
    ap.add_argument("--cpu-steps", type=int, default=0, help="steady-state iterations of the CPU step-loop sample (0 = pick by workload)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-cpu-loop", action="store_true", help="reference arm: skip the whole-loop difference of two solves")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle's check of the end-to-end solution")
    ap.add_argument("--replicas", action="store_true", help="N > 1: independent replicas instead of one column-sharded problem")
    ap.add_argument("--watchdog-sec", type=int, default=1700, help="kill this process if the run takes longer (0 = off)")
    args = ap.parse_args()
    if args.watchdog_sec > 0:
        # Default SIGALRM action on purpose: it ends the process even inside a native call that never
        # returns (a collective waiting for a rank that is gone), where a Python handler would not run.
        signal.alarm(args.watchdog_sec)

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    name = args.workload
    threads = os.cpu_count() or 1

    if args.impl == "reference":
        if rank == 0:
            reference_arm(args, name, world)
        return

    import torch
    import torch.distributed as dist

    import pdlp_b200

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("cpu:gloo,cuda:nccl", device_id=torch.device("cuda", local_rank))
    sharded = world > 1 and not args.replicas
    shard = starts = whole_qp = sharing = None
    num_shared = m_global = None
    if sharded:
        qp, starts, num_shared, m_global, whole_qp, sharing = make_block(name, args.scale, rank, world)
        unique_id = pdlp_b200.sharding.broadcast_unique_id()
        shard = pdlp_b200.ColumnShard(rank, world, int(starts[-1]), int(starts[rank]), unique_id, num_shared)
    else:
        qp = make_problem(name, args.scale, seed=0)
    K = qp.constraint_matrix
    m, n = K.shape
    nnz = K.nnz
    is_qp = qp.objective_matrix is not None
    b_step = bytes_per_step(n, m, nnz, is_qp, sharded)  # this rank's
    n_global, nnz_global, b_step_all = n, nnz, b_step
    if sharded:
        tot = torch.tensor([nnz, b_step], dtype=torch.float64, device="cuda")
        dist.all_reduce(tot)
        n_global, nnz_global, b_step_all = int(starts[-1]), int(tot[0].item()), float(tot[1].item())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- resident kernel-path timing ----
    dp = pdlp_b200.DeviceProblem(qp, device=local_rank, shard=shard)
    dp.rescale(5, True)  # the solver iterates on the rescaled problem
    stats = dp.stats()
    step0 = 1.0 / max(1e-20, stats.constraint_matrix_abs_max)
    w0 = stats.objective_vector_l2_norm / stats.combined_bounds_l2_norm if stats.objective_vector_l2_norm > 0 and stats.combined_bounds_l2_norm > 0 else 1.0
    x, y, st, _, _ = dp.trial_steps(args.warmup, step0, w0)
    barrier()
    with ClockSampler(local_rank) as clocks:
        # keep sampling for at least ~0.5 s so the clocks line has data
        t_begin = time.time()
        x2, y2, st2, seconds, launches = dp.trial_steps(args.steps, st["step_size"], w0, x, y)
        barrier()
        # Sharded trial steps are collective: every rank must repeat them the same number of times, so
        # the count comes from rank 0's clock, not from each rank's own.
        extra = torch.tensor([max(0, min(50, math.ceil((0.6 - (time.time() - t_begin)) / max(seconds, 1e-3))))], dtype=torch.int64, device="cuda")
        if world > 1:
            dist.broadcast(extra, src=0)
        extra_runs = int(extra.item())
        for _ in range(extra_runs):
            dp.trial_steps(args.steps, st["step_size"], w0, x, y)
    # Every enqueued trial must have run: a trial that forces numerical termination turns the rest of
    # the batch into no-op launches, which would still count towards steps/s.  (Rejected trials of an
    # iteration that has not accepted a step yet are in neither iterations_completed nor
    # num_rejected_steps: count executed trials, not their sum.)
    if st2["total_trials"] != args.steps or st2["forced_numerical_termination"]:
        raise SystemExit(f"bench: {st2['total_trials']} of {args.steps} timed trial steps executed ({st2['iterations_completed']} accepted, {st2['num_rejected_steps']} + "
                         f"{st2['pending_rejected_steps']} rejected; forced numerical termination {st2['forced_numerical_termination']}: movement {st2['last_movement']}, "
                         f"nonlinearity {st2['last_nonlinearity']}, step size {st2['step_size']}); no line printed")
    t = torch.tensor([seconds], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    seconds_max = float(t.item())
    steps_per_sec = (1 if sharded else world) * args.steps / seconds_max
    # per-kernel breakdown (second pass with CUDA events around every launch)
    breakdown = dp.trial_steps_profiled(args.steps, st["step_size"], w0, x, y)
    kb = kernel_bytes(n, m, nnz, is_qp)
    peak, peak_kind = measured_peak()
    kernels = []
    for kname, sec in breakdown.items():
        kernels.append({"kernel": kname, "us_per_launch": sec * 1e6, "algorithmic_bytes": kb[kname],
                        "achieved_gbs": kb[kname] / sec / 1e9 if sec > 0 and kb[kname] > 0 else None, "share": None})
    total_k = sum(k["us_per_launch"] for k in kernels)
    for k in kernels:
        k["share"] = k["us_per_launch"] / total_k if total_k > 0 else None
    dominant = max((k for k in kernels if k["algorithmic_bytes"] > 0), key=lambda k: k["us_per_launch"])
    step_gbs = b_step_all * args.steps / seconds_max / 1e9  # all ranks' algorithmic bytes over the slowest rank's time
    if not sharded:
        step_gbs = b_step * args.steps / seconds / 1e9
    peak_all = peak * (world if sharded else 1)

    dp.close()  # the resident copy goes away before the end-to-end solve uploads its own

    # ---- end-to-end solve from host buffers ----
    e2e = None
    parity = None
    solve_info = {}
    if not args.no_e2e:
        p = solve_params(name)
        barrier()
        t0 = time.time()
        if sharded:
            # Same communicator as the resident problem above: NCCL initialisation and the CUDA IPC mappings of the peer
            # exchange are per-process set-up (like the CUDA context), paid by the first sharded call and cached.
            result = pdlp_b200.primal_dual_hybrid_gradient_sharded(qp, p, device=local_rank, starts=starts, gather=False, num_shared_constraints=num_shared,
                                                                   unique_id=unique_id)
        else:
            result = pdlp_b200.primal_dual_hybrid_gradient(qp, p, device=local_rank)
        torch.cuda.synchronize()
        wall = time.time() - t0
        tw = torch.tensor([wall], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tw, op=dist.ReduceOp.MAX)
        wall_max = float(tw.item())
        its = result.solve_log.iteration_count
        counters = pdlp_b200.last_solve_counters()
        h2d = 16 * nnz + 8 * (n + 1) + 8 * (3 * n + 2 * m) + (8 * n if is_qp else 0)
        d2h = 8 * (2 * n + m)
        ci = [c for c in result.solve_log.solution_stats.convergence_information if c.candidate_type == result.solve_log.solution_type]
        ci = ci or list(result.solve_log.solution_stats.convergence_information)
        e2e = {"value": (1 if sharded else world) * its / wall_max, "unit": "iterations/s", "h2d_bytes_per_step": h2d / max(1, its), "d2h_bytes_per_step": d2h / max(1, its),
               "time_to_tolerance_sec": wall_max, "iterations": its, "termination_reason": int(result.solve_log.termination_reason),
               "h2d_bytes_per_solve": h2d, "d2h_bytes_per_solve": d2h}
        loop_sec = counters.get("loop_seconds") or 0.0
        trials = counters.get("trial_steps") or 0
        solve_info = {"solve_iterations": its, "solve_trial_steps": trials, "solve_wall_sec": wall_max,
                      "solve_setup_sec": counters.get("setup_seconds"), "solve_loop_sec": loop_sec,
                      "solve_device_loop_sec": counters.get("device_loop_seconds"), "solve_kernel_launches": counters.get("kernel_launches"),
                      "solve_collectives": counters.get("collectives"), "solve_upload_sec": counters.get("upload_seconds"), "solve_nccl_symmetric_buffers": counters.get("symmetric_buffers"), "solve_peer_exchanges": counters.get("peer_exchanges"),
                      "primal_objective": ci[0].primal_objective if len(ci) else None,
                      "relative_kkt_eps": workload_eps(name),
                      # the solver's own iteration rate (termination checks and restarts inside) and the
                      # roofline fraction of that loop: the trial steps' algorithmic bytes over the loop's wall time.
                      "loop_iterations_per_sec": its / loop_sec if loop_sec > 0 else None,
                      "loop_roofline_frac": b_step_all * trials / loop_sec / 1e9 / peak_all if loop_sec > 0 else None}
        if not args.no_parity:
            # Outside every timed region.  Sharded: the whole solution is gathered on every rank, rank 0 checks it.
            xs, ys = result.primal_solution, result.dual_solution
            check_qp = qp
            if sharded and whole_qp is not None:
                parts = [None] * world
                dist.all_gather_object(parts, (xs, ys))
                xs = np.concatenate([q[0] for q in parts])
                ys = sharing.gather([q[1] for q in parts])
                check_qp = whole_qp
            structure_check = name == "c5" and (sharded or args.scale > 2) and len(ci) > 0
            if structure_check:
                # Collective: every rank checks its own block; small vectors are added over the ranks.
                def reduce_sum(v):
                    if world == 1:
                        return v
                    t = torch.from_numpy(np.ascontiguousarray(v, dtype=np.float64)).cuda()
                    dist.all_reduce(t)
                    return t.cpu().numpy()
                t_check = time.time()
                parity = transportation_parity(qp, xs, ys, sharded, ci[0].primal_objective, its, reduce_sum)
                parity["ok"] = bool(compare_with_n1(parity, name, args.scale, world, its, parity["checker_primal_objective"], parity["eps"]) and parity["ok"])
                parity["seconds"] = time.time() - t_check
            elif rank == 0 and (not sharded or whole_qp is not None) and len(ci):
                parity = parity_check(check_qp, name, args.scale, xs, ys, ci[0].primal_objective, its, world)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        res = cpu_step_leg(qp, name, threads, args.cpu_steps or CPU_STEPS[name])
        cpu = {"value": res["value"], "unit": "iterations/s", "cores": threads, "kind": "port",
               "sample": f"{res['trial_steps']} steady-state trial steps ({res['accepted_steps']} accepted) of the CPU oracle's step loop after 2 untimed ones on the same workload "
                         f"(reference's threaded Sharder structure, num_shards = 4 x {threads}; no termination checks or restarts in the sample: the unit of `value`); "
                         f"{res['seconds']:.2f} s, preprocessing {res['preprocessing_seconds']:.2f} s excluded"}

    if rank == 0:
        line = {
            "metric": "pdhg_iterations_per_sec", "value": steps_per_sec, "unit": "iterations/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * seconds_max / args.steps, "higher_is_better": True, "scaling": "strong" if sharded or world == 1 else "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": make_config(name, args.scale, n_global, m_global if sharded else m, nnz_global, world, sharded),
            "value_unit_note": "trial steps (accepted + rejected) per second of the resident problem, no termination checks; loop_iterations_per_sec is the solver's own iteration rate",
            "clocks": clocks.summary(), "e2e": e2e, "gpu_launches": int(launches), "parity": parity,
            "nvlink": {"exchange": "peer memory (CUDA IPC): every rank stores its partial sums of the shared rows into every peer, then sums its N slots locally in rank order"
                                   if solve_info.get("solve_peer_exchanges") else "ncclAllReduce",
                       "pushed_bytes_per_step_per_rank": (world - 1) * (8 * num_shared + 8 * 4), "shared_constraints": num_shared, "constraints_on_rank0": m} if sharded else None,
            "roofline": {"bound": "hbm", "achieved": step_gbs, "peak": peak_all, "unit": "GB/s",
                         "frac": step_gbs / peak_all,
                         "traffic": None if sharded or args.scale != 1.0 else recorded_traffic(name, dominant["kernel"]),
                         "traffic_note": "dram bytes read + written per launch of the dominant kernel, from the committed ncu --set full capture (profiles/traffic.json); compare with that kernel's algorithmic_bytes",
                         "peak_kind": peak_kind + (f" x {world} GPUs" if sharded else ""), "unit_of_work": "one PDHG trial step (4 kernels: primal step + x average, K x + dual step + y average, K^T y + step test, decision)",
                         "algorithmic_bytes_per_step": b_step_all,
                         "dominant_kernel": dominant["kernel"], "dominant_kernel_achieved": dominant["achieved_gbs"],
                         "dominant_kernel_frac": (dominant["achieved_gbs"] or 0) / peak, "kernels": kernels},
            "cpu_baseline": cpu,
            "accepted_steps_in_timed_region": st2["iterations_completed"], "rejected_steps_in_timed_region": st2["num_rejected_steps"] + st2["pending_rejected_steps"],
        }
        line.update(solve_info)
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

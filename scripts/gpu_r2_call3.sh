#!/bin/bash
# Round 2, call 3 (one GPU): the no-alloc entry point's test alone in a fresh process, then the whole GPU suite
# (the same test last), then a full default bench line (parity object, loop rate, CPU step baseline) on C3 and C2.
mkdir -p gpurun_out
( timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q ) > gpurun_out/no_alloc_alone.log 2>&1; echo "alone exit $?" | tee -a gpurun_out/no_alloc_alone.log
tail -5 gpurun_out/no_alloc_alone.log
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu3.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu3.log
tail -4 gpurun_out/pytest_gpu3.log
timeout 900 python bench.py --steps 200 --warmup 5 > gpurun_out/bench_c3_r2.json 2> gpurun_out/bench_c3_r2.err; echo "bench c3 exit $?"
timeout 600 python bench.py --workload c2 --steps 200 --warmup 5 > gpurun_out/bench_c2_r2.json 2> gpurun_out/bench_c2_r2.err; echo "bench c2 exit $?"
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_c3_ref_r2.json 2> gpurun_out/bench_c3_ref_r2.err; echo "ref exit $?"
python - <<'P'
import json
for f in ("bench_c3_r2", "bench_c2_r2", "bench_c3_ref_r2"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
        print(f, {k: d.get(k) for k in ("value", "ms_per_step", "loop_iterations_per_sec", "loop_roofline_frac", "solve_loop_sec", "solve_iterations")}, "e2e", (d.get("e2e") or {}).get("value"),
              "frac", (d.get("roofline") or {}).get("frac"), "parity", d.get("parity"), "cpu", d.get("cpu_baseline"))
    except Exception as e:
        print(f, "failed", e)
P

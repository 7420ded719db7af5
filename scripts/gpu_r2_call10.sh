#!/bin/bash
# Round 2, call 10 (one GPU): the GPU tests that call 9 stopped in front of (the presolved entry point had no ctypes
# argtypes), then the full 1e9-nnz C5 LP on ONE GPU with the end-to-end solve: the N = 1 leg of the strong-scaling table.
mkdir -p gpurun_out
free -g | head -2
( time timeout 900 python -m pytest tests/test_presolve.py tests/test_logging.py tests/test_solve_cli.py tests/test_cpp_header.py tests/test_gpu_trial_steps.py \
    tests/test_zz_no_alloc.py tests/test_gpu_solve.py -m gpu -x -q ) > gpurun_out/pytest_gpu10.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu10.log
tail -4 gpurun_out/pytest_gpu10.log
( time env PDLP_B200_DEBUG_TIMING=1 timeout 1700 python bench.py --workload c5 --scale 8 --steps 100 --warmup 5 --no-cpu --watchdog-sec 1600 > gpurun_out/r2c_c5_s8_1gpu.json ) 2> gpurun_out/r2c_c5_s8_1gpu.err; echo "bench c5 s8 exit $?"
grep "loop phases\|^real" gpurun_out/r2c_c5_s8_1gpu.err; tail -3 gpurun_out/r2c_c5_s8_1gpu.err
cp /tmp/pdlp_b200_bench_n1_c5_s8.json gpurun_out/ 2>/dev/null
python - <<'P'
import json
for f in ("r2c_c5_s8_1gpu",):
    try:
        d = json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith("{")][-1])
        print(f, {k: d.get(k) for k in ("value", "ms_per_step", "loop_iterations_per_sec", "loop_roofline_frac", "solve_loop_sec", "solve_iterations", "solve_trial_steps", "solve_setup_sec", "solve_upload_sec", "solve_wall_sec", "primal_objective")},
              "e2e", d.get("e2e"), "frac", (d.get("roofline") or {}).get("frac"), "parity", d.get("parity"))
    except Exception as e:
        print(f, "failed", e)
P

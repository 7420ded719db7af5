#!/bin/bash
# Round 2, call 9 (one GPU): the GPU tests added since call 3, the loop's phase breakdown on C3 / C2
# (PDLP_B200_DEBUG_TIMING), then C5 on one GPU: scale 1 and the full 1e9-nnz LP (scale 8), both with the end-to-end solve.
mkdir -p gpurun_out
free -g | head -2; nproc
( time timeout 900 python -m pytest tests/test_presolve.py tests/test_logging.py tests/test_solve_cli.py tests/test_cpp_header.py tests/test_gpu_trial_steps.py \
    tests/test_zz_no_alloc.py tests/test_gpu_solve.py -m gpu -x -q ) > gpurun_out/pytest_gpu9.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu9.log
tail -4 gpurun_out/pytest_gpu9.log
( timeout 600 python -m pytest tests -m gpu -x -q -k "max_norm or localized" ) > gpurun_out/pytest_gpu9_bounds.log 2>&1; echo "bounds pytest exit $?"; tail -2 gpurun_out/pytest_gpu9_bounds.log
PDLP_B200_DEBUG_TIMING=1 timeout 900 python bench.py --steps 200 --warmup 5 --no-cpu > gpurun_out/r2c_c3_1gpu.json 2> gpurun_out/r2c_c3_1gpu.err; echo "bench c3 exit $?"
grep "loop phases" gpurun_out/r2c_c3_1gpu.err
PDLP_B200_DEBUG_TIMING=1 timeout 600 python bench.py --workload c2 --steps 200 --warmup 5 --no-cpu > gpurun_out/r2c_c2_1gpu.json 2> gpurun_out/r2c_c2_1gpu.err; echo "bench c2 exit $?"
grep "loop phases" gpurun_out/r2c_c2_1gpu.err
PDLP_B200_DEBUG_TIMING=1 timeout 900 python bench.py --workload c5 --scale 1 --steps 100 --warmup 5 --no-cpu > gpurun_out/r2c_c5_s1_1gpu.json 2> gpurun_out/r2c_c5_s1_1gpu.err; echo "bench c5 s1 exit $?"
grep "loop phases" gpurun_out/r2c_c5_s1_1gpu.err
( /usr/bin/time -v env PDLP_B200_DEBUG_TIMING=1 timeout 1700 python bench.py --workload c5 --scale 8 --steps 100 --warmup 5 --no-cpu --watchdog-sec 1600 > gpurun_out/r2c_c5_s8_1gpu.json ) 2> gpurun_out/r2c_c5_s8_1gpu.err; echo "bench c5 s8 exit $?"
grep "loop phases\|Maximum resident\|Elapsed" gpurun_out/r2c_c5_s8_1gpu.err
python - <<'P'
import json
for f in ("r2c_c3_1gpu", "r2c_c2_1gpu", "r2c_c5_s1_1gpu", "r2c_c5_s8_1gpu"):
    try:
        d = json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith("{")][-1])
        print(f, {k: d.get(k) for k in ("value", "ms_per_step", "loop_iterations_per_sec", "loop_roofline_frac", "solve_loop_sec", "solve_iterations", "solve_trial_steps", "solve_setup_sec", "solve_upload_sec", "solve_wall_sec")},
              "e2e", d.get("e2e"), "frac", (d.get("roofline") or {}).get("frac"), "parity", (d.get("parity") or {}).get("ok"))
    except Exception as e:
        print(f, "failed", e)
P

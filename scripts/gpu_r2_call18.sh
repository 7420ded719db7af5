#!/bin/bash
# Round 2, call 18 (one GPU): host events with one products pass (K x of current / average / difference and K^T y of
# average / difference from one stream of each matrix copy; KKT statistics and restart bounds as dense passes on them).
# The whole GPU suite, smoke, then the C3 / C2 / C5 lines with the loop's phase breakdown.
mkdir -p gpurun_out
( time timeout 1700 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu18.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu18.log
tail -5 gpurun_out/pytest_gpu18.log
( timeout 600 python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/smoke18.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/smoke18.log
for w in c3 c2 c5; do
  PDLP_B200_DEBUG_TIMING=1 timeout 900 python bench.py --workload $w --steps 200 --warmup 5 --no-cpu > gpurun_out/r2f_${w}_1gpu.json 2> gpurun_out/r2f_${w}_1gpu.err; echo "bench $w exit $?"
  grep "loop phases" gpurun_out/r2f_${w}_1gpu.err
done
python - <<'P'
import json
for f in ("r2f_c3_1gpu", "r2f_c2_1gpu", "r2f_c5_1gpu"):
    try:
        d = json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith("{")][-1])
        print(f, {k: d.get(k) for k in ("value", "ms_per_step", "loop_iterations_per_sec", "loop_roofline_frac", "solve_loop_sec", "solve_iterations", "solve_trial_steps", "solve_setup_sec", "solve_upload_sec", "solve_wall_sec")},
              "e2e", (d.get("e2e") or {}).get("value"), "frac", (d.get("roofline") or {}).get("frac"), "parity", (d.get("parity") or {}).get("ok"))
    except Exception as e:
        print(f, "failed", e)
P

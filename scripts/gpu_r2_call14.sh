#!/bin/bash
# Round 2, call 14 (one GPU): the whole GPU suite on the tree with the device-side interrupt, the fused restart bounds
# and the presolve hand-off; smoke; the C3 / C2 lines with the loop's phase breakdown.
mkdir -p gpurun_out
( time timeout 1700 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu14.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu14.log
tail -5 gpurun_out/pytest_gpu14.log
( timeout 600 python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/smoke14.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/smoke14.log
PDLP_B200_DEBUG_TIMING=1 timeout 900 python bench.py --steps 200 --warmup 5 > gpurun_out/r2d_c3_1gpu.json 2> gpurun_out/r2d_c3_1gpu.err; echo "bench c3 exit $?"
grep "loop phases" gpurun_out/r2d_c3_1gpu.err
PDLP_B200_DEBUG_TIMING=1 timeout 600 python bench.py --workload c2 --steps 200 --warmup 5 --no-cpu > gpurun_out/r2d_c2_1gpu.json 2> gpurun_out/r2d_c2_1gpu.err; echo "bench c2 exit $?"
grep "loop phases" gpurun_out/r2d_c2_1gpu.err
python - <<'P'
import json
for f in ("r2d_c3_1gpu", "r2d_c2_1gpu"):
    try:
        d = json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith("{")][-1])
        print(f, {k: d.get(k) for k in ("value", "ms_per_step", "loop_iterations_per_sec", "loop_roofline_frac", "solve_loop_sec", "solve_iterations", "solve_trial_steps", "solve_setup_sec", "solve_upload_sec", "solve_wall_sec")},
              "e2e", (d.get("e2e") or {}).get("value"), "frac", (d.get("roofline") or {}).get("frac"), "parity", (d.get("parity") or {}).get("ok"), "cpu", (d.get("cpu_baseline") or {}).get("value"))
    except Exception as e:
        print(f, "failed", e)
P

#!/bin/bash
# Round 2, call 25 (one GPU), the last one: the whole GPU suite on the final tree (matrix scaling over work blocks), then
# the default bench line (C3 with cpu_baseline, parity, e2e, loop phases).
mkdir -p gpurun_out
( time timeout 600 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu25.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu25.log
tail -4 gpurun_out/pytest_gpu25.log
PDLP_B200_DEBUG_TIMING=1 timeout 300 python bench.py > gpurun_out/r2k_c3_1gpu.json 2> gpurun_out/r2k_c3_1gpu.err; echo "bench exit $?"
grep "loop phases" gpurun_out/r2k_c3_1gpu.err
python - <<'P'
import json
try:
    d = json.loads([l for l in open("gpurun_out/r2k_c3_1gpu.json") if l.startswith("{")][-1])
    print({k: d.get(k) for k in ("value", "ms_per_step", "loop_iterations_per_sec", "loop_roofline_frac", "solve_loop_sec", "solve_setup_sec", "solve_upload_sec", "solve_wall_sec")},
          "e2e", (d.get("e2e") or {}).get("value"), "frac", d["roofline"]["frac"], "parity", (d.get("parity") or {}).get("ok"), "cpu", (d.get("cpu_baseline") or {}).get("value"))
except Exception as e:
    print("failed", e)
P

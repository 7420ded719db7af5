#!/bin/bash
# Round 2, call 1 (one GPU): trial-step unit check, A/B of the lean row-stream variants, device iteration-count distribution.
mkdir -p gpurun_out
( timeout 300 python scripts/check_trial_steps_gpu.py ) > gpurun_out/check_trial_steps.log 2>&1; echo "trial-steps check exit $?" | tee -a gpurun_out/check_trial_steps.log
for w in c3 c2 c4; do
  for v in base lean lean4; do
    lib=$PWD/or-tools_b200/lib/libpdlp_b200.so
    [ "$v" != base ] && lib=$PWD/or-tools_b200/lib_$v/libpdlp_b200.so
    [ -f "$lib" ] || continue
    PDLP_B200_LIBRARY=$lib timeout 300 python bench.py --workload $w --steps 100 --warmup 5 --no-cpu --no-e2e > gpurun_out/lean_${v}_${w}.json 2> gpurun_out/lean_${v}_${w}.err
    python - <<P
import json
try:
    d = json.load(open("gpurun_out/lean_${v}_${w}.json"))
    print("${v} ${w}", round(d["ms_per_step"], 4), "ms", round(d["roofline"]["frac"], 3), [(k["kernel"], round(k["us_per_launch"], 1)) for k in d["roofline"]["kernels"]])
except Exception as e:
    print("${v} ${w} failed", e)
P
  done
done
( time timeout 900 python scripts/iteration_distribution.py device gpurun_out/device_400.json --seeds 20 --perturb 6 ) > gpurun_out/device_400.log 2>&1
tail -3 gpurun_out/device_400.log

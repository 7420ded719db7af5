#!/bin/bash
# Round 2, call 4 (one GPU): A/B of the register-pipelined row-stream kernel (lib_pipe3: 3 CTAs/SM, lib_pipe4: 4) against the default.
mkdir -p gpurun_out
for w in c3 c2 c4 c5; do
  for v in base pipe3 pipe4; do
    lib=$PWD/or-tools_b200/lib/libpdlp_b200.so
    [ "$v" != base ] && lib=$PWD/or-tools_b200/lib_$v/libpdlp_b200.so
    [ -f "$lib" ] || continue
    PDLP_B200_LIBRARY=$lib timeout 400 python bench.py --workload $w --steps 100 --warmup 5 --no-cpu --no-e2e > gpurun_out/ab4_${v}_${w}.json 2> gpurun_out/ab4_${v}_${w}.err
    python - <<P
import json
try:
    d = json.load(open("gpurun_out/ab4_${v}_${w}.json"))
    print("${v} ${w}", round(d["ms_per_step"], 4), "ms", round(d["roofline"]["frac"], 3), [(k["kernel"], round(k["us_per_launch"], 1)) for k in d["roofline"]["kernels"]])
except Exception as e:
    print("${v} ${w} failed", e)
P
  done
done
PDLP_B200_LIBRARY=$PWD/or-tools_b200/lib_pipe4/libpdlp_b200.so timeout 600 python -m pytest tests/test_gpu_solve.py tests/test_gpu_trial_steps.py -m gpu -x -q > gpurun_out/pytest_pipe4.log 2>&1; echo "pytest pipe4 exit $?"; tail -2 gpurun_out/pytest_pipe4.log

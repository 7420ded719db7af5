#!/bin/bash
# Round 2, call 2 (one GPU): A/B of the asynchronous row-stream pipeline (default build) against the round-1
# kernel (lib_plain, -DPDLP_ROWSTREAM_ASYNC=0) and a 1024-nonzero / 4-CTA variant (lib_c1024); the GPU suite on
# the default build; smoke; one ncu --set full capture of the two async kernels.
mkdir -p gpurun_out
for w in c3 c2 c4 c5; do
  for v in base plain c1024; do
    lib=$PWD/or-tools_b200/lib/libpdlp_b200.so
    [ "$v" != base ] && lib=$PWD/or-tools_b200/lib_$v/libpdlp_b200.so
    [ -f "$lib" ] || continue
    PDLP_B200_LIBRARY=$lib timeout 400 python bench.py --workload $w --steps 100 --warmup 5 --no-cpu --no-e2e > gpurun_out/ab_${v}_${w}.json 2> gpurun_out/ab_${v}_${w}.err
    python - <<P
import json
try:
    d = json.load(open("gpurun_out/ab_${v}_${w}.json"))
    print("${v} ${w}", round(d["ms_per_step"], 4), "ms", round(d["roofline"]["frac"], 3), [(k["kernel"], round(k["us_per_launch"], 1)) for k in d["roofline"]["kernels"]])
except Exception as e:
    print("${v} ${w} failed", e)
P
  done
done
( time timeout 1500 python -m pytest tests -m gpu -x -q -s ) > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -n "typical_ratio\|passed\|failed\|pytest exit" gpurun_out/pytest_gpu.log | tail -12
( timeout 300 python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log; tail -4 gpurun_out/smoke.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:rowstream_async -c 4 -o gpurun_out/r02_async_c3 python bench.py --workload c3 --steps 2 --warmup 1 --no-cpu --no-e2e > gpurun_out/ncu_async.log 2>&1
echo "ncu exit $?"

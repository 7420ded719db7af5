#!/bin/bash
# One GPU: validation of the single index buffer (GPU suite) and kernels-only lines for it and
# for the 3-CTA variant.
mkdir -p gpurun_out
for v in base mE; do
  lib=or-tools_b200/lib/libpdlp_b200.so
  [ "$v" != base ] && lib=or-tools_b200/lib_$v/libpdlp_b200.so
  for w in c3 c2; do
    PDLP_B200_LIBRARY=$PWD/$lib timeout 300 python bench.py --workload $w --steps 100 --warmup 5 --no-cpu --no-e2e > gpurun_out/v9_${v}_${w}.json 2> gpurun_out/v9_${v}_${w}.err
  done
done
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu8.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu8.log
ls -la gpurun_out | tail -8

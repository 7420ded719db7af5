#!/bin/bash
# One GPU: A/B of the lean row-stream staging path (PDLP_ROWSTREAM_LEAN=1, pdlp_skeletons.cuh).
# Build the variant HERE first (the box only runs prebuilt files):
#   make -C or-tools_b200/csrc OUT=../lib_lean OBJ=../lib_lean/obj EXTRA_NVCCFLAGS=-DPDLP_ROWSTREAM_LEAN=1
#   make -C or-tools_b200/csrc OUT=../lib_lean4 OBJ=../lib_lean4/obj EXTRA_NVCCFLAGS="-DPDLP_ROWSTREAM_LEAN=1 -DPDLP_ROWSTREAM_REDUCE_UNROLL=4"
# then: scripts/gpurun_retry.sh --timeout 1500 -- bash scripts/gpu_call_lean.sh
# The variant must pass the whole GPU suite before its bench lines mean anything.
mkdir -p gpurun_out
( timeout 300 python scripts/check_trial_steps_gpu.py ) > gpurun_out/check_trial_steps.log 2>&1; echo "trial-steps check exit $?" | tee -a gpurun_out/check_trial_steps.log
lean=$PWD/or-tools_b200/lib_lean/libpdlp_b200.so
[ -f "$lean" ] || { echo "missing $lean" | tee gpurun_out/lean_missing.txt; exit 1; }
( time PDLP_B200_LIBRARY=$lean timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu_lean.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu_lean.log
tail -3 gpurun_out/pytest_gpu_lean.log
for v in base lean lean4; do
  lib=$PWD/or-tools_b200/lib/libpdlp_b200.so
  [ "$v" != base ] && lib=$PWD/or-tools_b200/lib_$v/libpdlp_b200.so
  [ -f "$lib" ] || continue
  for w in c3 c2 c4; do
    PDLP_B200_LIBRARY=$lib timeout 300 python bench.py --workload $w --steps 100 --warmup 5 --no-cpu --no-e2e > gpurun_out/lean_${v}_${w}.json 2> gpurun_out/lean_${v}_${w}.err
    tail -c 300 gpurun_out/lean_${v}_${w}.json
  done
done

#!/bin/bash
# compute-sanitizer over the small fixtures ("race detection / sanitizers"): memcheck, racecheck, synccheck
# and initcheck on the unit-level GPU tests that touch every kernel family once (row-stream, map-reduce, step decision,
# KKT passes, restart bounds, rescaling, matrix build).  Prepared in round 2 after the GPU budget was spent; NOT RUN YET.
# The CPU stand-ins that did run are the kernel emulation with guard pages behind every array and the ThreadSanitizer
# check of the shared-memory protocol.
#   gpurun --timeout 1500 -- bash scripts/gpu_sanitizer.sh
set -u
mkdir -p gpurun_out
TESTS="tests/test_reference_units.py tests/test_gpu_trial_steps.py"
for tool in memcheck racecheck synccheck initcheck; do
  # --target-processes all: pytest imports the library in-process, but the C++ caller tests fork.
  timeout 1200 compute-sanitizer --tool $tool --target-processes all --error-exitcode 97 --log-file gpurun_out/sanitizer_$tool.log \
    python -m pytest $TESTS -x -q -m gpu > gpurun_out/sanitizer_${tool}_pytest.log 2>&1
  echo "$tool: exit $?" | tee -a gpurun_out/sanitizer_summary.txt
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY" gpurun_out/sanitizer_$tool.log | tail -2 | tee -a gpurun_out/sanitizer_summary.txt
done

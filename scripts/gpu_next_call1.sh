#!/bin/bash
# First GPU call of the next round (one GPU; prepared at the end of round 2, when the budget was spent):
#   1. the whole GPU suite - 14 device tests were added after the last device run (test_python_surface.py,
#      the C++ / C samples, tests/test_zy_late_additions.py: no-row / no-column problems, the with-presolve small-value
#      checks, two host threads on one GPU);
#   2. smoke() and the default bench line;
#   3. compute-sanitizer on the unit-level tests (scripts/gpu_sanitizer.sh).
#   gpurun --timeout 2400 -- bash scripts/gpu_next_call1.sh
mkdir -p gpurun_out
( time timeout 600 python -m pytest tests -m gpu -q ) > gpurun_out/pytest_gpu_next1.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu_next1.log
grep -n "passed\|failed\|error" gpurun_out/pytest_gpu_next1.log | tail -3
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_next1.log 2>&1; echo "smoke exit $?"
timeout 900 python bench.py > gpurun_out/bench_next1.json 2> gpurun_out/bench_next1.err; echo "bench exit $?"
bash scripts/gpu_sanitizer.sh

#!/bin/bash
# Round 2, call 24 (one GPU): the whole GPU suite and smoke on the final tree, then the profiling passes of
# scripts/gpu_r2_ncu.sh (ncu launch list of the bench command, one `ncu --set full` capture of every distinct kernel).
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu24.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu24.log
tail -4 gpurun_out/pytest_gpu24.log
( timeout 300 python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/smoke24.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/smoke24.log
bash scripts/gpu_r2_ncu.sh

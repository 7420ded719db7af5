#!/bin/bash
# Round 2, call 15 (one GPU): the whole GPU suite (call 14 stopped at the interrupt test, whose flag went up while the
# process was still creating its CUDA context).
mkdir -p gpurun_out
( time timeout 1700 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu15.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu15.log
tail -5 gpurun_out/pytest_gpu15.log

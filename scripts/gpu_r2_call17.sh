#!/bin/bash
# Round 2, call 17 (one GPU): the whole GPU suite (calls 14 and 15 stopped at the interrupt test: first its flag went up while the
# process was still creating its CUDA context, then the host kept queueing the rest of its batch behind the stopped loop).
mkdir -p gpurun_out
( time timeout 1700 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu17.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu17.log
tail -5 gpurun_out/pytest_gpu17.log

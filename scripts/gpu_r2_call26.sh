#!/bin/bash
# Round 2, call 26 (one GPU): the GPU suite on the shipped binary (NVTX ranges, checked CUDA calls at the end of Solve).
mkdir -p gpurun_out
( time timeout 300 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu26.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu26.log
tail -4 gpurun_out/pytest_gpu26.log

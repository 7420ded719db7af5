#!/bin/bash
# Round 2, call 28 (one GPU, the last seconds of the budget): the presolve hand-off tests, with the two reference tests
# PresolveSolvesToOptimality and PresolveInfeasible in hand-off form.
mkdir -p gpurun_out
( timeout 45 python -m pytest tests/test_presolve.py -m gpu -q ) > gpurun_out/pytest_gpu28.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu28.log
tail -25 gpurun_out/pytest_gpu28.log

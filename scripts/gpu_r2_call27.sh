#!/bin/bash
# Round 2, call 27 (one GPU, the last seconds of the budget): the feasibility-polishing behaviour tests on the device
# with the time assertions of FeasibilityPolishingWorkStatsAreCorrect added.
mkdir -p gpurun_out
( timeout 45 python -m pytest tests/test_reference_pdhg_behaviour.py -m gpu -x -q -k polishing ) > gpurun_out/pytest_gpu27.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu27.log
tail -3 gpurun_out/pytest_gpu27.log

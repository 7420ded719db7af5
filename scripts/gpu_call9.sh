#!/bin/bash
# One GPU: validation of the multi-candidate trust-region root find (GPU suite), then the round's
# final 1-GPU lines: smoke, full bench (C3 with the CPU leg, C2), reference arm, ncu launch list.
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu9.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu9.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke9.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke9.log
( time timeout 600 python bench.py ) > gpurun_out/bench_c3_final.json 2> gpurun_out/bench_c3_final.err
timeout 300 python bench.py --workload c2 > gpurun_out/bench_c2_final.json 2> gpurun_out/bench_c2_final.err
( time timeout 600 python bench.py --impl reference --steps 8 --warmup 1 ) > gpurun_out/bench_c3_reference_final.json 2> gpurun_out/bench_c3_reference_final.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches_c3_final.csv python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_launches_final.log 2>&1
ls -la gpurun_out | tail -10

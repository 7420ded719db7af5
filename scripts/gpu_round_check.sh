#!/bin/bash
# One-GPU round check: parity tests, smoke, bench lines (ours + reference arm), ncu launch list and
# one ncu --set full capture of every kernel family.  Everything lands in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks_throttle_reasons.active --format=csv > gpurun_out/nvsmi.txt 2>&1
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
( time timeout 600 python bench.py ) > gpurun_out/bench_c3_full.json 2> gpurun_out/bench_c3_full.err
( time timeout 600 python bench.py --impl reference --steps 8 --warmup 1 ) > gpurun_out/bench_c3_reference.json 2> gpurun_out/bench_c3_reference.err
timeout 300 python bench.py --workload c2 > gpurun_out/bench_c2_full.json 2> gpurun_out/bench_c2_full.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches_c3.csv python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pdlp_b200 -c 140 -f -o gpurun_out/prof_all_c3 python scripts/profile_kernels.py c3 > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out

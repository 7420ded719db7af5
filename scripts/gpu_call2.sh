#!/bin/bash
mkdir -p gpurun_out
timeout 300 ./scripts/microbench/gather_bench > gpurun_out/gather_bench.csv 2>&1
( time timeout 900 python -m pytest tests/test_mp_model.py tests/test_gpu_solve.py tests/test_cabi.py -m gpu -x -q ) > gpurun_out/pytest_gpu2.log 2>&1
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke2.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke2.log
( time timeout 600 python bench.py --no-cpu ) > gpurun_out/bench_c3_stager.json 2> gpurun_out/bench_c3_stager.err
timeout 300 python bench.py --workload c2 --no-cpu > gpurun_out/bench_c2_stager.json 2> gpurun_out/bench_c2_stager.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'rowstream_kernel|map_reduce_kernel|decide_adaptive_kernel|finalize_kernel|scale_matrix_kernel' -c 160 -f -o gpurun_out/prof_all_c3 python scripts/profile_kernels.py c3 > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out

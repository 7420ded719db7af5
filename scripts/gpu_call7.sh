#!/bin/bash
# One GPU: validation of the per-block bank pad (GPU suite), kernels-only lines, C5 scale 1 (the
# 1-GPU leg of the weak-scaling comparison), and the upload timing with / without the overlapped partition.
mkdir -p gpurun_out
for w in c3 c2; do
  timeout 300 python bench.py --workload $w --steps 100 --warmup 5 --no-cpu --no-e2e > gpurun_out/v8_kernels_$w.json 2> gpurun_out/v8_kernels_$w.err
done
timeout 400 python bench.py --workload c5 --scale 1 --steps 50 --no-cpu --no-e2e > gpurun_out/bench_c5_1gpu.json 2> gpurun_out/bench_c5_1gpu.err
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu7.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu7.log
for i in 1 2; do
  timeout 300 python bench.py --workload c2 --steps 20 --no-cpu > gpurun_out/e2e_async_c2_$i.json 2>> gpurun_out/e2e.err
  PDLP_B200_SYNC_PARTITION=1 timeout 300 python bench.py --workload c2 --steps 20 --no-cpu > gpurun_out/e2e_sync_c2_$i.json 2>> gpurun_out/e2e.err
done
timeout 300 python bench.py --steps 20 --no-cpu > gpurun_out/e2e_async_c3.json 2>> gpurun_out/e2e.err
PDLP_B200_SYNC_PARTITION=1 timeout 300 python bench.py --steps 20 --no-cpu > gpurun_out/e2e_sync_c3.json 2>> gpurun_out/e2e.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:rowstream_kernel -s 20 -c 4 -f -o gpurun_out/prof_rowstream_c3_v8 python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_rowstream8.log 2>&1
ncu -i gpurun_out/prof_rowstream_c3_v8.ncu-rep --page raw --csv > gpurun_out/prof_rowstream_c3_v8.raw.csv 2> /dev/null
rm -f gpurun_out/prof_rowstream_c3_v8.ncu-rep
ls -la gpurun_out | tail -20

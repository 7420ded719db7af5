#!/bin/bash
# Two-GPU box: the whole GPU suite (single-GPU parity + the sharded tests), 1-GPU and 2-GPU bench
# lines, and the ncu capture of the row-stream kernels of the current tree.
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt 2>&1
( time timeout 1800 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu4.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu4.log
( time timeout 600 python bench.py ) > gpurun_out/bench_c3_v7.json 2> gpurun_out/bench_c3_v7.err
timeout 300 python bench.py --workload c2 --no-cpu > gpurun_out/bench_c2_v7.json 2> gpurun_out/bench_c2_v7.err
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 ) > gpurun_out/bench_c3_v7_2gpu.json 2> gpurun_out/bench_c3_v7_2gpu.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:rowstream_kernel -s 20 -c 4 -f -o gpurun_out/prof_rowstream_c3_v7 python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_rowstream7.log 2>&1
timeout 600 ncu --set full --clock-control none -k regex:map_reduce_kernel -s 40 -c 2 -f -o gpurun_out/prof_map_c3_v7 python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_map7.log 2>&1
for r in prof_rowstream_c3_v7 prof_map_c3_v7; do
  if [ -f gpurun_out/$r.ncu-rep ]; then
    ncu -i gpurun_out/$r.ncu-rep --page raw --csv > gpurun_out/$r.raw.csv 2> /dev/null
    sz=$(stat -c %s gpurun_out/$r.ncu-rep)
    if [ "$sz" -gt 20000000 ]; then rm -f gpurun_out/$r.ncu-rep; fi
  fi
done
rm -f gpurun_out/prof_*_v6.ncu-rep
du -sh gpurun_out; ls -la gpurun_out | tail -25

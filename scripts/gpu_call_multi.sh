#!/bin/bash
# usage: gpu_call_multi.sh N   (N = 2, 4 or 8 GPUs of one box)
# N = 2: the sharded parity tests + the C3 bench line.  Any N: C3 (strong scaling, the default
# workload) and C5 at scale N (the transportation LP with N x 1.25e8 nonzeros: weak scaling against
# the 1-GPU scale-1 line).
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus_$N.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
if [ "$N" = 2 ]; then
  ( time timeout 1200 python -m pytest tests/test_gpu_sharded.py -m gpu -x -q ) > gpurun_out/pytest_sharded.log 2>&1
  echo "pytest exit $?" >> gpurun_out/pytest_sharded.log
fi
( time timeout 900 $TR --master-port 29517 bench.py --gpus $N --no-cpu ) > gpurun_out/bench_c3_${N}gpu.json 2> gpurun_out/bench_c3_${N}gpu.err
( time timeout 900 $TR --master-port 29518 bench.py --gpus $N --workload c5 --scale $N --steps 50 --no-cpu --no-e2e ) > gpurun_out/bench_c5_${N}gpu.json 2> gpurun_out/bench_c5_${N}gpu.err
tail -c 600 gpurun_out/bench_c3_${N}gpu.err gpurun_out/bench_c5_${N}gpu.err
ls -la gpurun_out | tail -12

#!/bin/bash
# Second GPU call of the next round (8 GPUs: gpurun --gpus 8): the shipped peer path has run at N = 2 only since the
# finishing kernel's polling and the scalar exchange were parallelised (commit "Peer exchange: one polling thread per
# CTA ..."); the N = 4 / 8 lines in profiles/r02 are one commit older.  Sharded tests at 2 GPUs, then C3 at 2 / 4 / 8.
#   gpurun --gpus 8 --timeout 1500 -- bash scripts/gpu_next_call2.sh
mkdir -p gpurun_out
( time timeout 600 python -m pytest tests/test_gpu_sharded.py -m gpu -q ) > gpurun_out/pytest_sharded_next2.log 2>&1; echo "sharded exit $?"
for n in 2 4 8; do
  PDLP_B200_DEBUG_TIMING=1 timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2956$n \
    bench.py --gpus $n --workload c3 --steps 200 --no-cpu > gpurun_out/next2_c3_${n}gpu.json 2> gpurun_out/next2_c3_${n}gpu.err
  echo "N=$n exit $?"
done

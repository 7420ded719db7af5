#!/bin/bash
# usage: gpu_r2_multi.sh N [tests]   (N = 2, 4 or 8 GPUs of one box; "tests" also runs the sharded GPU tests, needs N >= 2)
# C3 (strong scaling) in both exchange modes (peer memory, and PDLP_B200_PEER_EXCHANGE=0 = ncclAllReduce) and C5 at
# scale 8 (the 1e9-nnz LP, strong scaling) in peer mode.
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus_$N.txt 2>&1
nvidia-smi topo -m > gpurun_out/topo_$N.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
if [ "$2" = tests ]; then
  ( time timeout 1200 python -m pytest tests/test_gpu_sharded.py -m gpu -x -q -s ) > gpurun_out/pytest_sharded_peer.log 2>&1; echo "pytest peer exit $?" | tee -a gpurun_out/pytest_sharded_peer.log
  grep -n "passed\|failed\|error" gpurun_out/pytest_sharded_peer.log | tail -3
  ( time PDLP_B200_PEER_EXCHANGE=0 timeout 1200 python -m pytest tests/test_gpu_sharded.py -m gpu -x -q ) > gpurun_out/pytest_sharded_nccl.log 2>&1; echo "pytest nccl exit $?" | tee -a gpurun_out/pytest_sharded_nccl.log
  grep -n "passed\|failed\|error" gpurun_out/pytest_sharded_nccl.log | tail -3
fi
show() {
python - "$1" <<'P'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
    print(sys.argv[1], "value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 4), "frac", round(d["roofline"]["frac"], 3),
          [(k["kernel"], round(k["us_per_launch"], 1)) for k in d["roofline"]["kernels"]], "e2e", d.get("e2e"), "peer", d.get("solve_peer_exchanges"),
          "collectives", d.get("solve_collectives"), "parity", (d.get("parity") or {}).get("ok"), (d.get("parity") or {}).get("vs_n1"))
except Exception as e:
    print(sys.argv[1], "failed", e)
P
}
if [ "$N" = 1 ]; then
  timeout 900 python bench.py --gpus 1 --no-cpu --steps 200 > gpurun_out/r2_c3_1gpu.json 2> gpurun_out/r2_c3_1gpu.err; show gpurun_out/r2_c3_1gpu.json
else
  timeout 900 $TR --master-port 29517 bench.py --gpus $N --no-cpu --steps 200 > gpurun_out/r2_c3_${N}gpu_peer.json 2> gpurun_out/r2_c3_${N}gpu_peer.err; show gpurun_out/r2_c3_${N}gpu_peer.json
  PDLP_B200_PEER_EXCHANGE=0 timeout 900 $TR --master-port 29518 bench.py --gpus $N --no-cpu --steps 200 --no-parity > gpurun_out/r2_c3_${N}gpu_nccl.json 2> gpurun_out/r2_c3_${N}gpu_nccl.err; show gpurun_out/r2_c3_${N}gpu_nccl.json
fi
tail -c 400 gpurun_out/r2_c3_${N}gpu*.err

#!/bin/bash
# Round 2, call 23 (two GPUs): the sharded GPU tests on the tree with the split fused peer pass and its in-kernel signal,
# in peer mode (finish kernel: one polling thread per CTA; scalar exchange and flag stores by one thread per peer), then C3 at N = 2 (kernels only).
mkdir -p gpurun_out
( time timeout 600 python -m pytest tests/test_gpu_sharded.py -m gpu -x -q ) > gpurun_out/pytest_sharded23.log 2>&1; echo "pytest sharded (peer) exit $?" | tee -a gpurun_out/pytest_sharded23.log
grep -n "passed\|failed\|error" gpurun_out/pytest_sharded23.log | tail -2
PDLP_B200_DEBUG_TIMING=1 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 2 --workload c3 --steps 200 --no-cpu --no-e2e > gpurun_out/r2j_c3_2gpu.json 2> gpurun_out/r2j_c3_2gpu.err
python - <<'P'
import json
try:
    d = json.loads([l for l in open("gpurun_out/r2j_c3_2gpu.json") if l.startswith("{")][-1])
    print("N=2", "ms/step", round(d["ms_per_step"], 4), [(k["kernel"], round(k["us_per_launch"], 1)) for k in d["roofline"]["kernels"]])
except Exception as e:
    print("failed", e)
P
grep "K x pass in peer mode" gpurun_out/r2j_c3_2gpu.err | sort | head -2

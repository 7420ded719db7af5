#!/bin/bash
# 2 GPUs: does the early stop of the C5 trial-step batch depend on the exchange mode or on the problem size?
mkdir -p gpurun_out
run() {  # tag scale steps env...
  local tag=$1 scale=$2 steps=$3; shift 3
  env "$@" timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --workload c5 --scale $scale --steps $steps --no-cpu --no-e2e > gpurun_out/c5dbg_$tag.json 2> gpurun_out/c5dbg_$tag.err
  echo "$tag: exit $?"; grep "^bench:" gpurun_out/c5dbg_$tag.err | head -1
  python - gpurun_out/c5dbg_$tag.json <<'P'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
    print("  ms/step", round(d["ms_per_step"], 4), "accepted", d["accepted_steps_in_timed_region"], "rejected", d["rejected_steps_in_timed_region"], [(k["kernel"], round(k["us_per_launch"], 1)) for k in d["roofline"]["kernels"]])
except Exception as e:
    print("  no line:", e)
P
}
run s2_peer_100 2 100 A=1
run s2_nccl_100 2 100 PDLP_B200_PEER_EXCHANGE=0
run s8_peer_100 8 100 A=1
run s8_nccl_100 8 100 PDLP_B200_PEER_EXCHANGE=0

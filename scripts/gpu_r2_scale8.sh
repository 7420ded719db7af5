#!/bin/bash
# Round 2 (8-GPU box): C3 strong scaling with the peer-memory exchange at N = 8 and 4, and the NCCL path at N = 8 for the A/B.
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_8.txt 2>&1
export PDLP_B200_DEBUG_TIMING=1
show() {
python - "$1" <<'P'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
    print(sys.argv[1], "value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 4), "frac", round(d["roofline"]["frac"], 3),
          [(k["kernel"], round(k["us_per_launch"], 1)) for k in d["roofline"]["kernels"]], "e2e", round(d["e2e"]["value"], 1), d["e2e"]["time_to_tolerance_sec"], d["e2e"]["iterations"],
          "loop", d.get("loop_iterations_per_sec"), "peer", d.get("solve_peer_exchanges"), "parity", (d.get("parity") or {}).get("ok"), (d.get("parity") or {}).get("vs_n1"))
except Exception as e:
    print(sys.argv[1], "failed", e)
P
}
run() {  # N mode port
  local N=$1 mode=$2 port=$3
  local TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port"
  if [ "$mode" = nccl ]; then export PDLP_B200_PEER_EXCHANGE=0; else unset PDLP_B200_PEER_EXCHANGE; fi
  timeout 600 $TR bench.py --gpus $N --no-cpu --steps 200 > gpurun_out/r2_c3_${N}gpu_${mode}.json 2> gpurun_out/r2_c3_${N}gpu_${mode}.err
  show gpurun_out/r2_c3_${N}gpu_${mode}.json
  grep "peer set-up" gpurun_out/r2_c3_${N}gpu_${mode}.err | head -3
}
run 8 peer 29521
run 8 nccl 29522
run 4 peer 29523

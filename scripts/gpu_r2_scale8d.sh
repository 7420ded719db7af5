#!/bin/bash
# Round 2 (8-GPU box), fourth scaling call: the fused peer pass with a static CTA split (shared rows on a share of the
# grid, private rows on the rest, concurrently).  C3 at N = 8, kernels only: the heuristic split, no split, 35 %.
mkdir -p gpurun_out
show() {
python - "$1" <<'P'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
    print(sys.argv[1], "value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 4), "frac", round(d["roofline"]["frac"], 3),
          [(k["kernel"], round(k["us_per_launch"], 1)) for k in d["roofline"]["kernels"]])
except Exception as e:
    print(sys.argv[1], "failed", e)
P
}
run() {  # tag port env...
  local tag=$1 port=$2; shift 2
  env PDLP_B200_DEBUG_TIMING=1 "$@" timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port bench.py --gpus 8 --workload c3 --steps 200 \
      --no-cpu --no-e2e > gpurun_out/r2g_c3_8gpu_$tag.json 2> gpurun_out/r2g_c3_8gpu_$tag.err
  show gpurun_out/r2g_c3_8gpu_$tag.json
  grep "K x pass in peer mode" gpurun_out/r2g_c3_8gpu_$tag.err | sort | head -2
  grep "^bench:\|Error" gpurun_out/r2g_c3_8gpu_$tag.err | head -3
}
run auto 29541 A=1
run nosplit 29542 PDLP_B200_PEER_SHARED_CTAS_PCT=0
run pct35 29543 PDLP_B200_PEER_SHARED_CTAS_PCT=35

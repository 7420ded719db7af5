#!/bin/bash
# Round 2 (8-GPU box), third scaling call: the CTA-wide peer push.  Sharded GPU tests (2 ranks, incl. the random
# projections), then C3 strong scaling at N = 8, 4, 2 (kernels + end-to-end solve; vs_n1 from the committed N = 1 line).
mkdir -p gpurun_out
python - <<'P'
import json
d = json.loads([l for l in open("profiles/r02/r2d_c3_1gpu.json") if l.startswith("{")][-1])
json.dump({"iterations": d["parity"]["iterations"], "primal_objective": d["parity"]["oracle_primal_objective"]}, open("/tmp/pdlp_b200_bench_n1_c3_s1.json", "w"))
P
( time timeout 900 python -m pytest tests/test_gpu_sharded.py -m gpu -x -q ) > gpurun_out/pytest_sharded_push.log 2>&1; echo "pytest sharded exit $?" | tee -a gpurun_out/pytest_sharded_push.log
grep -n "passed\|failed\|error" gpurun_out/pytest_sharded_push.log | tail -2
show() {
python - "$1" <<'P'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
    e = d.get("e2e") or {}
    print(sys.argv[1], "value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 4), "frac", round(d["roofline"]["frac"], 3),
          [(k["kernel"], round(k["us_per_launch"], 1)) for k in d["roofline"]["kernels"]], "e2e", e.get("value"), e.get("time_to_tolerance_sec"), e.get("iterations"), e.get("termination_reason"),
          "loop", d.get("loop_iterations_per_sec"), "peer", d.get("solve_peer_exchanges"), "parity", (d.get("parity") or {}).get("ok"), (d.get("parity") or {}).get("vs_n1"))
except Exception as e:
    print(sys.argv[1], "failed", e)
P
}
run() {  # N workload scale port extra...
  local N=$1 w=$2 scale=$3 port=$4; shift 4
  local TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port"
  PDLP_B200_DEBUG_TIMING=1 timeout 700 $TR bench.py --gpus $N --workload $w --scale $scale --no-cpu "$@" > gpurun_out/r2e_${w}_s${scale}_${N}gpu.json 2> gpurun_out/r2e_${w}_s${scale}_${N}gpu.err
  show gpurun_out/r2e_${w}_s${scale}_${N}gpu.json
  grep "K x pass in peer mode" gpurun_out/r2e_${w}_s${scale}_${N}gpu.err | sort | head -8
  grep "^bench:\|Error" gpurun_out/r2e_${w}_s${scale}_${N}gpu.err | head -3
}
run 8 c3 1.0 29531 --steps 200
run 4 c3 1.0 29532 --steps 200
run 2 c3 1.0 29533 --steps 200

#!/bin/bash
# Round 2 (8-GPU box), fifth scaling call: C3 at N = 8 on the final peer path (CTA split + in-kernel signal), full line
# with the end-to-end solve and the parity object (vs_n1 from the committed N = 1 line).
mkdir -p gpurun_out
python - <<'P'
import json
d = json.loads([l for l in open("profiles/r02/r2f_c3_1gpu.json") if l.startswith("{")][-1])
json.dump({"iterations": d["parity"]["iterations"], "primal_objective": d["parity"]["oracle_primal_objective"]}, open("/tmp/pdlp_b200_bench_n1_c3_s1.json", "w"))
P
PDLP_B200_DEBUG_TIMING=1 timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 8 --workload c3 --steps 200 --no-cpu \
    > gpurun_out/r2i_c3_8gpu.json 2> gpurun_out/r2i_c3_8gpu.err
python - <<'P'
import json
try:
    d = json.loads([l for l in open("gpurun_out/r2i_c3_8gpu.json") if l.startswith("{")][-1])
    e = d.get("e2e") or {}
    print("N=8", "value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 4), "frac", round(d["roofline"]["frac"], 3), [(k["kernel"], round(k["us_per_launch"], 1)) for k in d["roofline"]["kernels"]],
          "e2e", e.get("value"), e.get("time_to_tolerance_sec"), e.get("iterations"), "loop", d.get("loop_iterations_per_sec"), "parity", (d.get("parity") or {}).get("ok"), (d.get("parity") or {}).get("vs_n1"))
except Exception as e:
    print("failed", e)
P
grep "K x pass in peer mode" gpurun_out/r2i_c3_8gpu.err | sort | head -8
grep "^bench:\|Error" gpurun_out/r2i_c3_8gpu.err | head -3

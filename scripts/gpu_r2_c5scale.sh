#!/bin/bash
# C5 strong scaling: the full 1e9-nnz LP (scale 8) column-sharded over $1 GPUs, with the end-to-end solve and the
# structure-aware parity object; the N = 1 solve's (iterations, objective) of call 10 is installed for the vs_n1 entry.
N=${1:-8}
mkdir -p gpurun_out
cp profiles/r02/n1_reference_c5_s8.json /tmp/pdlp_b200_bench_n1_c5_s8.json
nvidia-smi topo -m > gpurun_out/topo_c5_$N.txt 2>&1
( time env PDLP_B200_DEBUG_TIMING=1 timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29547 bench.py --gpus $N --workload c5 --scale 8 \
    --steps 100 --warmup 5 --no-cpu --watchdog-sec 1400 > gpurun_out/r2c_c5_s8_${N}gpu.json ) 2> gpurun_out/r2c_c5_s8_${N}gpu.err; echo "bench c5 s8 N=$N exit $?"
grep "loop phases\|^real\|^bench:" gpurun_out/r2c_c5_s8_${N}gpu.err | head -12
python - $N <<'P'
import json, sys
f = f"r2c_c5_s8_{sys.argv[1]}gpu"
try:
    d = json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith("{")][-1])
    print(f, {k: d.get(k) for k in ("value", "ms_per_step", "loop_iterations_per_sec", "solve_loop_sec", "solve_iterations", "solve_trial_steps", "solve_setup_sec", "solve_upload_sec", "solve_wall_sec", "primal_objective")},
          "e2e", d.get("e2e"), "frac", (d.get("roofline") or {}).get("frac"), "parity", d.get("parity"), [(k["kernel"], round(k["us_per_launch"], 1)) for k in d["roofline"]["kernels"]])
except Exception as e:
    print(f, "failed", e)
P

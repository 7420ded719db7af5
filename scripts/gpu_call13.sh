#!/bin/bash
# One GPU: the config-4 line (diagonal-objective QP, 5e7 nnz, eps 1e-6).
mkdir -p gpurun_out
( time timeout 420 python bench.py --workload c4 --steps 100 ) > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err
tail -c 400 gpurun_out/bench_c4.err

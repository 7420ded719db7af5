#!/bin/bash
# One-GPU check of the current tree; every large artefact is summarised on the box and removed
# if it would push gpurun_out/ over the 64 MiB copy-back limit.
mkdir -p gpurun_out
timeout 300 ./scripts/microbench/gather_bench > gpurun_out/gather_bench.csv 2>&1
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu3.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu3.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke3.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke3.log
( time timeout 600 python bench.py --no-cpu ) > gpurun_out/bench_c3_v6.json 2> gpurun_out/bench_c3_v6.err
timeout 300 python bench.py --workload c2 --no-cpu > gpurun_out/bench_c2_v6.json 2> gpurun_out/bench_c2_v6.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:rowstream_kernel -s 20 -c 4 -f -o gpurun_out/prof_rowstream_c3_v6 python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_rowstream.log 2>&1
timeout 600 ncu --set full --clock-control none -k regex:map_reduce_kernel -s 40 -c 3 -f -o gpurun_out/prof_map_c3_v6 python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_map.log 2>&1
for r in prof_rowstream_c3_v6 prof_map_c3_v6; do
  if [ -f gpurun_out/$r.ncu-rep ]; then
    ncu -i gpurun_out/$r.ncu-rep --page raw --csv > gpurun_out/$r.raw.csv 2> /dev/null
    sz=$(stat -c %s gpurun_out/$r.ncu-rep)
    if [ "$sz" -gt 25000000 ]; then rm -f gpurun_out/$r.ncu-rep; echo "$r.ncu-rep removed ($sz bytes)" >> gpurun_out/ncu_rowstream.log; fi
  fi
done
du -sh gpurun_out; ls -la gpurun_out

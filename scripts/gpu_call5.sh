#!/bin/bash
# One GPU: kernel-variant comparison (kernels-only bench on C3 and C2), the GPU suite on the base
# library, full bench lines, ncu capture of the row-stream kernels.
mkdir -p gpurun_out
for v in base mA mB mC; do
  lib=or-tools_b200/lib/libpdlp_b200.so
  [ "$v" != base ] && lib=or-tools_b200/lib_$v/libpdlp_b200.so
  for w in c3 c2; do
    PDLP_B200_LIBRARY=$PWD/$lib timeout 300 python bench.py --workload $w --steps 100 --warmup 5 --no-cpu --no-e2e > gpurun_out/variant_${v}_${w}.json 2> gpurun_out/variant_${v}_${w}.err
  done
done
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu5.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu5.log
( time timeout 600 python bench.py ) > gpurun_out/bench_c3_v7.json 2> gpurun_out/bench_c3_v7.err
timeout 300 python bench.py --workload c2 > gpurun_out/bench_c2_v7.json 2> gpurun_out/bench_c2_v7.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:rowstream_kernel -s 20 -c 4 -f -o gpurun_out/prof_rowstream_c3_v7 python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_rowstream7.log 2>&1
timeout 600 ncu --set full --clock-control none -k regex:map_reduce_kernel -s 40 -c 2 -f -o gpurun_out/prof_map_c3_v7 python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_map7.log 2>&1
for r in prof_rowstream_c3_v7 prof_map_c3_v7; do
  if [ -f gpurun_out/$r.ncu-rep ]; then
    ncu -i gpurun_out/$r.ncu-rep --page raw --csv > gpurun_out/$r.raw.csv 2> /dev/null
    sz=$(stat -c %s gpurun_out/$r.ncu-rep)
    if [ "$sz" -gt 20000000 ]; then rm -f gpurun_out/$r.ncu-rep; fi
  fi
done
du -sh gpurun_out; ls -la gpurun_out | tail -30

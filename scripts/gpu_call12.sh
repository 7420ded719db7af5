#!/bin/bash
# One GPU, end of round: GPU suite on the final tree, then one `ncu --set full` capture of the FIRST
# launch of every distinct kernel (template instance) of a set-up + three trial steps + termination
# check + restart evaluation on C3 (scripts/profile_kernels.py); exported to CSV on the box.
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu12.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu12.log
timeout 120 python scripts/profile_kernels.py c3 > gpurun_out/profile_kernels_plain.log 2>&1
timeout 900 ncu --set full --clock-control none --kernel-name-base mangled --kernel-id :::1 -f -o gpurun_out/prof_every_kernel_c3 python scripts/profile_kernels.py c3 > gpurun_out/ncu_every_kernel.log 2>&1
ncu -i gpurun_out/prof_every_kernel_c3.ncu-rep --page raw --csv > gpurun_out/prof_every_kernel_c3.raw.csv 2> /dev/null
ls -la gpurun_out/prof_every_kernel_c3.* >> gpurun_out/ncu_every_kernel.log
rm -f gpurun_out/prof_every_kernel_c3.ncu-rep
( time timeout 300 python bench.py --no-cpu ) > gpurun_out/bench_c3_last.json 2> gpurun_out/bench_c3_last.err
ls -la gpurun_out | tail -8

#!/bin/bash
# Round 2, profiling call (one GPU): the ncu launch list of the bench command (per-launch durations; compare SHARES with
# the bench's own CUDA-event table) and one `ncu --set full` capture of the first launch of every distinct kernel.
mkdir -p gpurun_out
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2_bench_for_launches.json 2> gpurun_out/r2_bench_for_launches.err; echo "plain bench exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r02_launches_c3.csv python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_launches.log 2>&1; echo "ncu launch list exit $?"
timeout 300 python scripts/profile_kernels.py c3 > gpurun_out/profile_kernels_plain.log 2>&1; echo "plain profile script exit $?"; tail -2 gpurun_out/profile_kernels_plain.log
timeout 1500 ncu --set full --clock-control none --import-source on --kernel-name-base mangled --kernel-id :::1 -f -o gpurun_out/r02_every_kernel_c3 python scripts/profile_kernels.py c3 > gpurun_out/ncu_every_kernel.log 2>&1; echo "ncu set full exit $?"
ncu -i gpurun_out/r02_every_kernel_c3.ncu-rep --page raw --csv > gpurun_out/r02_every_kernel_c3.raw.csv 2> /dev/null
ls -la gpurun_out/r02_every_kernel_c3.* gpurun_out/r02_launches_c3.csv
rm -f gpurun_out/r02_every_kernel_c3.ncu-rep

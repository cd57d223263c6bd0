#!/bin/bash
# Round-2 profile pass (one GPU): launch list with DRAM bytes, then ncu --set full of the five kernels of the step.
set -x
B="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu --no-extras"
$B > gpurun_out/r02_prof_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/r02_launches.csv $B > gpurun_out/r02_ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:fft_tma_kernel --launch-skip 4 --launch-count 4 \
    -o gpurun_out/r02_fft -f $B > gpurun_out/r02_ncu_fft.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:merkle_leaf_tma_kernel --launch-skip 1 --launch-count 1 \
    -o gpurun_out/r02_leaf -f $B > gpurun_out/r02_ncu_leaf.log 2>&1
ls -la gpurun_out

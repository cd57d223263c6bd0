#!/bin/bash
set -x
export TSTWO_FFT_TC=3
B="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu --no-extras"
$B > gpurun_out/tc3_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:fft_tma_kernel --launch-skip 4 --launch-count 4 \
    -o gpurun_out/r02_fft_tc3 -f $B > gpurun_out/tc3_ncu.log 2>&1

#!/bin/bash
python scripts/pass_times.py 100 23 > gpurun_out/pass_times_23.txt 2>&1; cat gpurun_out/pass_times_23.txt
ncu --set full --clock-control none --import-source on -k regex:fft_tma_kernel --launch-skip 12 --launch-count 4 -o gpurun_out/r02_r9 -f python scripts/pass_times.py 100 23 > gpurun_out/r02_ncu_r9.log 2>&1
ls -la gpurun_out/r02_r9.ncu-rep

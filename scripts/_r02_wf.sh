#!/bin/bash
python scripts/widefib_run.py 20 | tail -3
ncu --set full --clock-control none --import-source on -k regex:widefib_mma_kernel --launch-skip 2 --launch-count 1 -o gpurun_out/r02_widefib -f python scripts/widefib_run.py 20 > gpurun_out/r02_ncu_wf.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:quotients_mma_kernel --launch-skip 2 --launch-count 1 -o gpurun_out/r02_quot -f python scripts/quot_run.py 22 > gpurun_out/r02_ncu_q.log 2>&1
ls -la gpurun_out/*.ncu-rep

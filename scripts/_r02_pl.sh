#!/bin/bash
python scripts/prove_launches.py 23
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/prove_launches.csv python scripts/prove_launches.py 23 > gpurun_out/prove_launches_ncu.txt 2>&1; tail -2 gpurun_out/prove_launches_ncu.txt

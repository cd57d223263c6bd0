#!/bin/bash
set -x
python -m pytest tests/test_gpu_fri_prover.py tests/test_gpu_poly.py tests/test_gpu_pcs.py tests/test_gpu_prove.py -x -q > gpurun_out/t_fri.log 2>&1; tail -3 gpurun_out/t_fri.log
python scripts/elementwise_bench.py > gpurun_out/elementwise.txt 2>&1
python scripts/prove_profile.py 23 100 > gpurun_out/prove_profile_23.txt 2>&1
python scripts/fri_commit_profile.py 24 > gpurun_out/fri_commit_plain.txt 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/fri_launches.csv python scripts/fri_commit_profile.py 24 > gpurun_out/fri_commit_ncu.txt 2>&1
python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu --no-extras > gpurun_out/b1.log 2>&1

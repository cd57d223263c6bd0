#!/bin/bash
set -x
python -m pytest tests/test_gpu_fri_prover.py tests/test_gpu_prove.py tests/test_gpu_pcs.py tests/test_gpu_merkle.py tests/test_gpu_commitment.py -x -q > gpurun_out/t_dec.log 2>&1; tail -5 gpurun_out/t_dec.log
python scripts/fri_bench.py 24 2>&1 | tail -3
python scripts/prove_profile.py 23 100 > gpurun_out/prove_profile_23b.txt 2>&1; head -30 gpurun_out/prove_profile_23b.txt

#!/bin/bash
set -x
python -m pytest tests/test_gpu_merkle.py tests/test_gpu_fri_prover.py tests/test_gpu_commitment.py tests/test_gpu_pipeline.py tests/test_gpu_prove.py -x -q > gpurun_out/t_quick.log 2>&1; tail -3 gpurun_out/t_quick.log
python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/b4.log 2> gpurun_out/b4.err; tail -c 200 gpurun_out/b4.err
python -c "
import json
d=json.loads(open('gpurun_out/b4.log').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['root'][:16], d['root_matches_oracle_checked_root'], d['extras']['config4_fri']['fri_commit_ms'], d['extras']['config5_prove']['prove_ms'], d['extras']['config5_prove']['trace_commit_ms'])
"

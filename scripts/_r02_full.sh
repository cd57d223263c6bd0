#!/bin/bash
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/t_full.log 2>&1; tail -4 gpurun_out/t_full.log
python scripts/fri_commit_profile.py 24 > gpurun_out/fri_commit_plain2.txt 2>&1; cat gpurun_out/fri_commit_plain2.txt
python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/b3.log 2> gpurun_out/b3.err; tail -c 200 gpurun_out/b3.err

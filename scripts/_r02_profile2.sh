#!/bin/bash
set -x
python -m pytest tests/test_gpu_env.py tests/test_zz_no_alloc.py -x -q > gpurun_out/t_env.log 2>&1; tail -3 gpurun_out/t_env.log
B="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu --no-extras"
$B > gpurun_out/r02_prof_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/r02_launches.csv $B > gpurun_out/r02_ncu_launch.log 2>&1
python scripts/fri_commit_profile.py 24 > gpurun_out/fri_commit_plain.txt 2>&1; cat gpurun_out/fri_commit_plain.txt
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/fri_launches.csv python scripts/fri_commit_profile.py 24 > gpurun_out/fri_commit_ncu.txt 2>&1
ncu --set full --clock-control none --import-source on -k regex:merkle_node_kernel --launch-skip 23 --launch-count 1 \
    -o gpurun_out/r02_node -f $B > gpurun_out/r02_ncu_node.log 2>&1

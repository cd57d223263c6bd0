#!/bin/bash
N=${1:-4}
set -x
timeout 300 python -m pytest tests/test_gpu_pipeline.py -x -q -k "overlapped or peer_flags" > gpurun_out/t_taper.log 2>&1; tail -3 gpurun_out/t_taper.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 scripts/overlap_sweep.py $SWEEP 2>&1 | grep -v "^W\|^\*\*\*" | tee gpurun_out/overlap_taper_n$N.txt

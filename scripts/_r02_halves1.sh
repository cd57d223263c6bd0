#!/bin/bash
# single-GPU checks of the chunked leaf hashing and the column-halves store path
timeout 900 python -m pytest tests/test_gpu_pipeline.py tests/test_gpu_peer_commit.py -x -q -k "absorb or halves or peer_sharded or overlapped or peer_flags" > gpurun_out/t_halves1.log 2>&1
echo "rc $?"; tail -15 gpurun_out/t_halves1.log

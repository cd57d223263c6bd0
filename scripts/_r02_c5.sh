#!/bin/bash
set -x
python -m pytest tests/test_gpu_prove.py tests/test_gpu_pipeline.py::test_peer_flags_signal_wait_put -x -q --durations=8 > gpurun_out/t_c5.log 2>&1; tail -15 gpurun_out/t_c5.log
python bench.py --steps 10 --warmup 3 > gpurun_out/b2.log 2> gpurun_out/b2.err; tail -c 300 gpurun_out/b2.err

set -x
mkdir -p gpurun_out
(time timeout 600 python -m pytest tests/test_gpu_pipeline.py -x -q -k "config3 or overlapped" --durations=3) 2>&1 | tail -12

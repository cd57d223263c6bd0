#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_fft_modes.py -x -q 2>&1 | tail -5
python scripts/pass_times.py 100 23

#!/bin/bash
python -m pytest tests/test_gpu_pcs.py tests/test_gpu_fri.py tests/test_gpu_prove.py tests/test_gpu_air.py -x -q 2>&1 | tail -2
python scripts/quot_run.py 22 | tail -3
python scripts/prove_launches.py 23

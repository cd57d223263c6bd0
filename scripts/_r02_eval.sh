#!/bin/bash
python -m pytest tests/test_gpu_poly.py tests/test_gpu_pcs.py tests/test_zz_no_alloc.py tests/test_gpu_prove.py -x -q 2>&1 | tail -3
python scripts/elementwise_bench.py 2>&1 | grep eval_at_point

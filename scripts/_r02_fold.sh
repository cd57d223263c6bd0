#!/bin/bash
python -m pytest tests/test_gpu_fri.py tests/test_gpu_fri_prover.py tests/test_gpu_air.py tests/test_gpu_config_sizes.py tests/test_gpu_prove.py tests/test_gpu_pcs.py -x -q 2>&1 | tail -3
python scripts/elementwise_bench.py 2>&1 | grep "accumulate\|decompose"
python scripts/fri_bench.py 24 2>&1 | grep -v "^$" | head -12

timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
timeout 300 python scripts/elementwise_bench.py 2>&1 | tail -12
timeout 300 python scripts/prove_bench.py 23 100 | head -2

timeout 300 python scripts/prove_bench.py 23 100
timeout 300 python scripts/prove_bench.py 23 100
timeout 100 python scripts/fri_bench.py 2>&1 | tail -3

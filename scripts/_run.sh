set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 400 python bench.py --no-cpu --no-e2e > gpurun_out/bench_try.json 2> gpurun_out/bench_try.err; python -c "
import json; d=json.load(open('gpurun_out/bench_try.json')); print(d['ms_per_step'], [(k, round(v['ms'],3)) for k,v in d['kernels'].items()])"

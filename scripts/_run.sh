set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 400 python bench.py --no-cpu > gpurun_out/bench_reorder.json 2> gpurun_out/bench_reorder.err; python -c "
import json; d=json.load(open('gpurun_out/bench_reorder.json')); print(d['ms_per_step'], d['kernels'], d['e2e']['ms_per_step'], d['roofline'].get('int_pipe'))"

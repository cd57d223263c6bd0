#!/bin/bash
python -m pytest tests -m gpu -x -q > gpurun_out/t_final.log 2>&1; tail -3 gpurun_out/t_final.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py --steps 10 --warmup 3 > gpurun_out/b_final.log 2> gpurun_out/b_final.err; tail -c 300 gpurun_out/b_final.err
python -c "
import json
d=json.loads(open('gpurun_out/b_final.log').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['root'][:16], d['root_matches_oracle_checked_root'], d['gpu_launches'])
print(json.dumps(d['extras']))
print(d['e2e']['ms_per_step'], d['cpu_baseline'])
print(d['roofline']['frac'], d['int_roofline']['fft']['frac'], d['int_roofline']['blake2s']['frac'])
"

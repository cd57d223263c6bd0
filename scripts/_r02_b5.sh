#!/bin/bash
python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/b5.log 2> gpurun_out/b5.err; tail -c 300 gpurun_out/b5.err
python -c "
import json
d=json.loads(open('gpurun_out/b5.log').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['root'][:16], d['root_matches_oracle_checked_root'])
print(json.dumps(d['extras'], indent=1))
print([ (r['kernel'][:40], r.get('traffic')) for r in d['rooflines']])
print(d['e2e'])
"

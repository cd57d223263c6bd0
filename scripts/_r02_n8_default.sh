#!/bin/bash
# the driver's N-GPU launch of our arm with default-ish arguments (e2e leg and phase breakdown included)
N=${1:-8}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/default_n$N.json 2> gpurun_out/default_n$N.err; echo "rc $?"
python - <<PY
import json
d = json.loads(open('gpurun_out/default_n$N.json').read().strip().splitlines()[-1])
print(d['n_gpus'], round(d['ms_per_step'], 4), d['root'][:16], d['root_matches_oracle_checked_root'], 'e2e', round(d['e2e']['ms_per_step'], 2), round(d['e2e']['e2e_over_floor'], 3), d['gpu_launches'])
print(d['column_ownership'][:40], d['roofline']['frac'], d['clocks'])
print({k: round(v['ms_max_over_ranks'], 3) for k, v in d['phases'].items() if isinstance(v, dict) and 'ms_max_over_ranks' in v})
PY

#!/bin/bash
# what the driver does at round end for one N: our arm and the reference arm under torchrun, default-ish arguments
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29521 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/scale_check_n$N.json 2> gpurun_out/scale_check_n$N.err; echo "rc $?"
timeout 600 $TR --master-port 29522 bench.py --impl reference --gpus $N --steps 1 --warmup 0 > gpurun_out/scale_check_ref_n$N.json 2> gpurun_out/scale_check_ref_n$N.err; echo "rc $?"
python - <<PY
import json
d = json.loads(open('gpurun_out/scale_check_n$N.json').read().strip().splitlines()[-1])
print(d['n_gpus'], round(d['ms_per_step'], 4), d['root'][:16], d['root_matches_oracle_checked_root'], 'e2e', round(d['e2e']['ms_per_step'], 2), round(d['e2e']['e2e_over_floor'], 3), d['gpu_launches'])
print(d['roofline']['frac'], d['clocks'])
r = json.loads(open('gpurun_out/scale_check_ref_n$N.json').read().strip().splitlines()[-1])
print('reference arm:', r['impl'], r['n_gpus'], round(r['ms_per_step'], 1), r['cpu_baseline']['cores'])
print(len(open('gpurun_out/scale_check_n$N.json').read().strip().splitlines()), 'line(s) on stdout (ours),', len(open('gpurun_out/scale_check_ref_n$N.json').read().strip().splitlines()), '(reference)')
PY

#!/bin/bash
N=${1:-2}
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n${N}b.log 2> gpurun_out/bench_n${N}b.err
python - <<PY
import json
for l in open('gpurun_out/bench_n${N}b.log').read().splitlines():
    try: d = json.loads(l)
    except Exception: continue
    print(d['n_gpus'], d['ms_per_step'], d['root'][:16], d['root_matches_oracle_checked_root'])
    print(json.dumps(d['roofline'])[:400])
    print(json.dumps(d['e2e']))
PY
tail -3 gpurun_out/bench_n${N}b.err

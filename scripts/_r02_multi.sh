#!/bin/bash
# usage: bash scripts/_r02_multi.sh N "sweep configs"   -- everything multi-GPU in one call (N ranks on one box)
N=${1:-8}
SWEEP=${2:-"1,0 4,62 4,58 4,66 2,62 8,62"}
set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/scale_n${N}_flags.log 2> gpurun_out/scale_n${N}_flags.err
timeout 300 $TR --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 --no-e2e --fence nccl > gpurun_out/scale_n${N}_nccl.log 2> gpurun_out/scale_n${N}_nccl.err
timeout 300 $TR --master-port 29514 bench.py --gpus $N --steps 5 --warmup 3 --e2e-chunks 8 > gpurun_out/scale_n${N}_e2e8.log 2> gpurun_out/scale_n${N}_e2e8.err
timeout 300 python bench.py --gpus $N --single-process --steps 20 --warmup 5 > gpurun_out/scale_n${N}_single.log 2> gpurun_out/scale_n${N}_single.err
timeout 400 $TR --master-port 29513 scripts/overlap_sweep.py $SWEEP 2>&1 | grep -v "^W\|^\*\*\*" > gpurun_out/overlap_n$N.txt
timeout 300 python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/t_multi_n$N.log 2>&1
for f in flags nccl e2e8 single; do python - <<PY
import json
for l in open('gpurun_out/scale_n${N}_$f.log').read().splitlines():
    try: d = json.loads(l)
    except Exception: continue
    print('$f', d.get('n_gpus'), 'ms', round(d['ms_per_step'], 4), d.get('root', '')[:16], 'e2e', (d.get('e2e') or {}).get('ms_per_step'), (d.get('e2e') or {}).get('h2d_only'))
    ph = d.get('phases') or {}
    print('   ', {k: (round(v['ms_max_over_ranks'], 3) if 'ms_max_over_ranks' in v else v.get('gbs_per_rank_over_the_lde_phase')) for k, v in ph.items() if isinstance(v, dict)})
PY
done
cat gpurun_out/overlap_n$N.txt; tail -2 gpurun_out/t_multi_n$N.log

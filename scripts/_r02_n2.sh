#!/bin/bash
# usage: bash scripts/_r02_n2.sh N   -- flags vs NCCL fences at N ranks, plus the one-process path
N=${1:-2}
set -x
timeout 300 python -m pytest tests/test_gpu_pipeline.py::test_peer_flags_signal_wait_put tests/test_gpu_multi.py -x -q > gpurun_out/t_n$N.log 2>&1; tail -3 gpurun_out/t_n$N.log
for fence in flags nccl flags; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --no-e2e --fence $fence > gpurun_out/bench_n${N}_$fence.log 2> gpurun_out/bench_n${N}_$fence.err
  tail -c 1500 gpurun_out/bench_n${N}_$fence.log | python -c "
import sys, json
for l in sys.stdin.read().splitlines():
    try: d = json.loads(l)
    except Exception: continue
    print('$fence', d['n_gpus'], 'ms', d['ms_per_step'], d['root'][:16], json.dumps(d.get('phases', {}))[:900])
"
done
timeout 300 python bench.py --gpus $N --single-process --steps 20 --warmup 5 > gpurun_out/bench_n${N}_single.log 2> gpurun_out/bench_n${N}_single.err; tail -c 600 gpurun_out/bench_n${N}_single.log

#!/bin/bash
# N GPUs: PeerShardedCommit parity under torchrun (both column ownerships), then the step time of each ownership / overlap setting
N=${1:-2}
shift
if [ -z "$SKIP_PYTEST" ]; then
  timeout 900 python -m pytest tests/test_gpu_peer_commit.py -x -q > gpurun_out/t_halves_n$N.log 2>&1
  echo "pytest rc $?"; tail -5 gpurun_out/t_halves_n$N.log
fi
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
port=29530
for cfg in "contiguous auto" "halves auto" "$@"; do
  set -- $cfg
  port=$((port + 1))
  timeout 300 $TR --master-port $port bench.py --gpus $N --steps 20 --warmup 3 --no-e2e --no-extras --ownership $1 --overlap $2 > gpurun_out/halves_n${N}_$1_$2.json 2> gpurun_out/halves_n${N}_$1_$2.err
  python - <<PY
import json
try:
    d = json.loads(open('gpurun_out/halves_n${N}_$1_$2.json').read().strip().splitlines()[-1])
    ph = d.get('phases', {})
    print('N=$N $1 $2: %.3f ms' % d['ms_per_step'], d['root'][:16], d['root_matches_oracle_checked_root'],
          ' | '.join('%s %.3f' % (k[:12], v['ms_max_over_ranks']) for k, v in ph.items() if isinstance(v, dict) and 'ms_max_over_ranks' in v))
except Exception as e:
    print('N=$N $1 $2 failed:', e); print(open('gpurun_out/halves_n${N}_$1_$2.err').read()[-1500:])
PY
done

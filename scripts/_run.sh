N=$1
for ex in "peer" "ce --chunks 2" "ce --chunks 4"; do
  echo "== $ex"
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 8 --warmup 3 --no-e2e --exchange $ex 2>&1 | grep '"metric"' | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['root'][:16])"
done

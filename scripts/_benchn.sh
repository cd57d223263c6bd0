mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 300 python bench.py --no-cpu --no-e2e --steps 5 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('N=1', d['ms_per_step'], d['clocks'])"
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
python -c "import sys,json; d=json.load(open('gpurun_out/bench_n$N.json')); print('N=$N', d['ms_per_step'], d['value'], d['clocks'], d.get('e2e'))"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 10 --warmup 3 --overlap off --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('N=$N overlap off', d['ms_per_step'], d['clocks'])"

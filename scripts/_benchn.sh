mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
nvidia-smi topo -m 2>/dev/null | head -14 > gpurun_out/topo_n$N.txt
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
python -c "import sys,json; d=json.load(open('gpurun_out/bench_n$N.json')); print('N=$N bound', d['ms_per_step'], d['clocks'], d.get('e2e'))"
TSTWO_NO_NUMA_BIND=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 10 --warmup 3 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('N=$N unbound', d['ms_per_step'], d.get('e2e'))"

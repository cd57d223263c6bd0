mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
python -c "import sys,json; d=json.load(open('gpurun_out/bench_n$N.json')); print('N=$N', d['ms_per_step'], d['value'], d['clocks'], d.get('e2e'), d['root'][:16])"
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus $N --steps 1 --warmup 1 2>/dev/null | cut -c1-300

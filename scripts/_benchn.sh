# both multi-GPU arms of bench.py on all GPUs of the box: torchrun (one process per GPU) and --single-process
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
python -c "import sys,json; d=json.load(open('gpurun_out/bench_n$N.json')); print('torchrun N=$N', d['ms_per_step'], d.get('e2e',{}).get('ms_per_step'), d['root'][:16]); print(json.dumps(d.get('phases')))" || tail -5 gpurun_out/bench_n$N.err
timeout 500 python bench.py --single-process --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_sp$N.json 2> gpurun_out/bench_sp$N.err
python -c "import sys,json; d=json.load(open('gpurun_out/bench_sp$N.json')); print('single-process N=$N', d['ms_per_step'], d.get('e2e',{}).get('ms_per_step'), d['root'][:16], d['gpu_launches'])" || tail -5 gpurun_out/bench_sp$N.err

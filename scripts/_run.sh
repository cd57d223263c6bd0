set -x
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 scripts/overlap_sweep.py 1,0 2,62 4,54 4,62 4,70 8,54 8,62 8,70 16,62 1,0 2>&1 | grep -v "^W\|^\*\*\*" | tee gpurun_out/overlap_n$N.txt | tail -14

set -x
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 scripts/overlap_sweep.py $SWEEP 2>&1 | grep -v "^W\|^\*\*\*" | tee gpurun_out/overlap_b_n$N.txt | tail -24

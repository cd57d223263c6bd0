mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
for cfg in "1 0" "4 62"; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 scripts/multi_phases.py $cfg 2>&1 | grep "N=\|ms" | tee -a gpurun_out/phases_n$N.txt
done

#!/bin/bash
python scripts/leaf_chunk_bench.py 20 22 > gpurun_out/leaf_chunk_bench.txt 2>&1; cat gpurun_out/leaf_chunk_bench.txt
ncu --set full --clock-control none --import-source on -k regex:merkle_leaf --launch-skip 12 --launch-count 4 -o gpurun_out/r02_chunk -f python scripts/leaf_chunk_bench.py 20 > gpurun_out/r02_ncu_chunk.log 2>&1
ls -la gpurun_out/r02_chunk.ncu-rep

set -x
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 400 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; tail -c 600 gpurun_out/bench_default.json
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 2>/dev/null | tail -1 | cut -c1-600
timeout 200 python scripts/elementwise_bench.py > gpurun_out/elementwise.txt 2>&1; tail -14 gpurun_out/elementwise.txt
timeout 200 python scripts/fri_bench.py > gpurun_out/fri_c4.txt 2>&1; tail -4 gpurun_out/fri_c4.txt
timeout 300 ncu --set full --clock-control none --import-source on -k regex:quotients_mma -c 1 -o gpurun_out/quot_mma python scripts/quot_run.py 22 2 > /dev/null 2>&1; ls -la gpurun_out/*.ncu-rep

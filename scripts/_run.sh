mkdir -p gpurun_out
timeout 300 python scripts/prove_profile.py 20 100 > gpurun_out/prove_profile.txt 2>&1; head -c 200 gpurun_out/prove_profile.txt

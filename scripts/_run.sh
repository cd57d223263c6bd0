timeout 600 python -m pytest tests/test_gpu_air.py -x -q 2>&1 | tail -4
timeout 120 python scripts/widefib_run.py
timeout 300 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:widefib_mma --csv python scripts/widefib_run.py 2>&1 | grep -i "widefib_" | awk -F'","' '{print substr($1,1,40), $(NF-2), $NF}' | tr -d '"' | sort | uniq -c | sort -k2 | head -4

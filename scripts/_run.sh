timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 120 python scripts/quot_modes.py
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none -k regex:quotients --csv python scripts/quot_modes.py 2>&1 | grep -i "quotients" | awk -F'","' '{print substr($1,1,40), $(NF-2), $NF}' | tr -d '"' | sort | uniq -c | sort -k2 | head -40
timeout 200 python scripts/fri_bench.py 2>&1 | tail -3

"""Per-launch durations / DRAM bytes from an `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv` log.
usage: python scripts/launch_list.py gpurun_out/launches.csv"""
import csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
h = rows[0]; ki = h.index("Kernel Name"); mi = h.index("Metric Name"); vi = h.index("Metric Value"); ii = h.index("ID")
d = {}
for r in rows[1:]:
    d.setdefault((int(r[ii]), r[ki][:64]), {})[r[mi]] = float(r[vi].replace(",", ""))
tot = 0.0
for (i, k), v in sorted(d.items()):
    ms = v.get("gpu__time_duration.sum", 0) / 1e6
    tot += ms
    gb = (v.get("dram__bytes_read.sum", 0) + v.get("dram__bytes_write.sum", 0)) / 1e9
    print(f"{i:3d} {k:64s} {ms:8.3f} ms  dram R {v.get('dram__bytes_read.sum', 0) / 1e9:6.2f} W {v.get('dram__bytes_write.sum', 0) / 1e9:6.2f} GB  {gb / ms if ms else 0:6.2f} TB/s")
print(f"total {tot:.3f} ms")

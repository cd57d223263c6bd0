"""Per-launch durations / DRAM bytes from an `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv` log.
usage: python scripts/launch_list.py gpurun_out/launches.csv [--json profiles/rNN_launches.json] [--skip N]

With --json it also writes the per-kernel means (launches, ms, DRAM bytes read / written per launch) that bench.py reads for
`roofline.traffic`: the file named in every roofline entry of the bench line is exactly that output."""
import csv, json, re, sys

args = sys.argv[1:]
out_json = args[args.index("--json") + 1] if "--json" in args else None
skip = int(args[args.index("--skip") + 1]) if "--skip" in args else 0
rows = [r for r in csv.reader(open(args[0])) if len(r) > 10]
h = rows[0]; ki = h.index("Kernel Name"); mi = h.index("Metric Name"); vi = h.index("Metric Value"); ii = h.index("ID")
d = {}
for r in rows[1:]:
    d.setdefault((int(r[ii]), r[ki]), {})[r[mi]] = float(r[vi].replace(",", ""))


def short(name):
    m = re.search(r"(\w+)(<[^>]*>)?\(", name)
    if not m:
        return name[:64]
    return m.group(1) + (m.group(2) or "").replace("(bool)", "").replace("(int)", "")


tot = 0.0
agg = {}
for n, ((i, k), v) in enumerate(sorted(d.items())):
    ms = v.get("gpu__time_duration.sum", 0) / 1e6
    tot += ms
    gb = (v.get("dram__bytes_read.sum", 0) + v.get("dram__bytes_write.sum", 0)) / 1e9
    print(f"{i:3d} {short(k):48s} {ms:8.3f} ms  dram R {v.get('dram__bytes_read.sum', 0) / 1e9:6.2f} W {v.get('dram__bytes_write.sum', 0) / 1e9:6.2f} GB  {gb / ms if ms else 0:6.2f} TB/s")
    if n >= skip:
        a = agg.setdefault(short(k), {"launches": 0, "ms": 0.0, "dram_read_bytes": 0.0, "dram_write_bytes": 0.0})
        a["launches"] += 1; a["ms"] += ms
        a["dram_read_bytes"] += v.get("dram__bytes_read.sum", 0); a["dram_write_bytes"] += v.get("dram__bytes_write.sum", 0)
print(f"total {tot:.3f} ms")
if out_json:
    for a in agg.values():
        for key in ("ms", "dram_read_bytes", "dram_write_bytes"):
            a[key] /= a["launches"]
    json.dump({"source": args[0], "note": "means per launch; ncu --clock-control none, cold-cache and serialised: compare shares, not absolutes",
               "kernels": agg}, open(out_json, "w"), indent=1, sort_keys=True)
    print("wrote", out_json)

"""Per-SASS-instruction view of one kernel of an .ncu-rep (source page): instruction mix, shared-memory wavefront
excess and the top stall sites.  usage: python scripts/ncu_sass_hot.py REP LAUNCH_INDEX [top]"""
import csv, io, subprocess, sys, collections
rep, idx = sys.argv[1], int(sys.argv[2])
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--launch-skip", str(idx),
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
print(rows[0][1])
hdr = rows[1]
H = {h: i for i, h in enumerate(hdr)}
data = rows[2:]
def f(r, k):
    try: return float(r[H[k]])
    except Exception: return 0.0
tot_inst = sum(f(r, "Instructions Executed") for r in data)
tot_samp = sum(f(r, "# Samples") for r in data)
mix = collections.Counter(); smp = collections.Counter()
for r in data:
    op = r[H["Source"]].split()
    op = [o for o in op if not o.startswith("@")]
    name = op[0] if op else "?"
    name = ".".join(name.split(".")[:2]) if name.startswith(("IMAD", "LDS", "STS", "LDG", "STG", "LEA", "LDGSTS")) else name.split(".")[0]
    mix[name] += f(r, "Instructions Executed"); smp[name] += f(r, "# Samples")
print(f"warp instructions {tot_inst:.0f}, samples {tot_samp:.0f}")
for k, v in mix.most_common(22):
    print(f"  {k:16s} {v / tot_inst * 100:6.2f} % inst   {smp[k] / tot_samp * 100:6.2f} % samples")
print("shared wavefronts: total %.0f ideal %.0f" % (sum(f(r, "L1 Wavefronts Shared") for r in data), sum(f(r, "L1 Wavefronts Shared Ideal") for r in data)))
ex = sorted(data, key=lambda r: -f(r, "L1 Wavefronts Shared Excessive"))[:8]
for r in ex:
    if f(r, "L1 Wavefronts Shared Excessive") > 0:
        print("  excess %10.0f of %10.0f  %s" % (f(r, "L1 Wavefronts Shared Excessive"), f(r, "L1 Wavefronts Shared"), r[H["Source"]][:90]))
print("L2 sectors global: total %.0f ideal %.0f" % (sum(f(r, "L2 Theoretical Sectors Global") for r in data), sum(f(r, "L2 Theoretical Sectors Global Ideal") for r in data)))
stall_cols = [h for h in hdr if h.startswith("stall_") and not h.endswith("(Not Issued)")]
print("top stall sites:")
for r in sorted(data, key=lambda r: -f(r, "# Samples"))[:top]:
    st = sorted(((f(r, c), c[6:]) for c in stall_cols), reverse=True)[:3]
    print("  %6.2f %%  %-70s %s" % (f(r, "# Samples") / tot_samp * 100, r[H["Source"]][:70], " ".join(f"{n}={v:.0f}" for v, n in st if v > 0)))

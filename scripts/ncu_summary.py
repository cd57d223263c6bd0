"""Summarise an .ncu-rep (ncu --set full) into the handful of per-kernel numbers the roofline discussion uses.
usage: python scripts/ncu_summary.py gpurun_out/prof.ncu-rep [--json profiles/rNN_name.json] > profiles/rNN_name.txt
--json also writes {kernel: counters} (last capture of each kernel wins), which bench.py cites in rooflines[].ncu."""
import csv, json, re, subprocess, sys, io
rep = sys.argv[1]
out_json = sys.argv[sys.argv.index("--json") + 1] if "--json" in sys.argv else None
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
want = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sectors_srcunit_tex_op_write.sum",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "memory_l1_wavefronts_shared", "memory_l1_wavefronts_shared_ideal"]
stalls = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio")]
print(f"# {rep}: ncu --set full --clock-control none (per launch; cold-cache, serialised)")
for r in data:
    print("-" * 100)
    for w in want:
        if w in hdr:
            i = hdr.index(w)
            print(f"{w:72s} {r[i]:>24s} {units[i]}")
    st = sorted(((float(r[hdr.index(h)]), h[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]) for h in stalls), reverse=True)
    print("warp stalls per issue: " + ", ".join(f"{n}={v:.2f}" for v, n in st if v >= 0.05))

if out_json:
    def short(name):
        m = re.search(r"(\w+)(<[^>]*>)?\(", name)
        return name[:64] if not m else m.group(1) + (m.group(2) or "").replace("(bool)", "").replace("(int)", "")
    def val(r, k):
        try:
            return float(r[hdr.index(k)].replace(",", ""))
        except Exception:
            return None
    MULT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    TM = {"ns": 1e-6, "nsecond": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "s": 1e3, "second": 1e3}
    def scaled(r, k, table):
        v = val(r, k)
        return None if v is None else v * table.get(units[hdr.index(k)], 1.0)
    ks = {}
    for r in data:
        ks[short(r[hdr.index("Kernel Name")])] = {
            "duration_ms": scaled(r, "gpu__time_duration.sum", TM),
            "registers_per_thread": val(r, "launch__registers_per_thread"),
            "issue_active_pct": val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
            "alu_pipe_active_pct": val(r, "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"),
            "fma_pipe_active_pct": val(r, "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
            "tensor_pipe_active_pct": val(r, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
            "lsu_wavefronts_pct": val(r, "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"),
            "dram_throughput_pct": val(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
            "dram_read_bytes": scaled(r, "dram__bytes_read.sum", MULT), "dram_write_bytes": scaled(r, "dram__bytes_write.sum", MULT),
            "warps_active_pct": val(r, "sm__warps_active.avg.pct_of_peak_sustained_active"),
        }
    json.dump({"source": rep, "note": "ncu --set full --clock-control none; one capture per kernel (the last one in the report)", "kernels": ks},
              open(out_json, "w"), indent=1, sort_keys=True)

"""Summarise an .ncu-rep (ncu --set full) into the handful of per-kernel numbers the roofline discussion uses.
usage: python scripts/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/rNN_name.txt"""
import csv, subprocess, sys, io
rep = sys.argv[1]
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

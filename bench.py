#!/usr/bin/env python
"""bench.py -- LDE + Blake2s-commit throughput (BASELINE.json metric) on N B200s of one node.

A step = one pass of the hot path over one batch of synthetic input: interpolate 256 trace columns of 2^22 M31
elements, evaluate them on the blow-up-2 canonic domain (2^23 rows), Blake2s Merkle commit, root on the host.
  value : trace elements / s with the trace resident in HBM (CUDA events, max over ranks)
  e2e   : the same through the C-ABI call a host makes (tstwo_lde_commit_host): trace in pinned HOST memory,
          host->device copies inside the timed region, root read back
N > 1 : columns are sharded by column for the LDE, an NCCL all-to-all re-shards by row range, every rank
        hashes its leaves and builds its subtree, rank 0 finishes the top log2(N) layers.
--impl reference : the CPU oracle (C restatement of TStwo's CpuBackend; the TypeScript reference cannot run here)
        on all host cores, on a bounded sample of the same workload.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

_REAL_STDOUT = sys.stdout
N_COLS, LOG_SIZE, LOG_BLOWUP = 256, 22, 1
# DRAM bytes (read + written) per launch of the kernels below at the headline size, from ncu (profiles/r01_launches_tma.txt)
NCU_TRAFFIC = {"merkle_leaf_layer": 8.59e9 + 0.27e9, "evaluate (2 fft launches)": (4.51e9 + 8.53e9 + 8.61e9 + 8.53e9) / 2,
               "interpolate (2 fft launches)": (4.31e9 + 4.25e9 + 4.31e9 + 4.23e9) / 2}
# (column groups, SMs of the NVLink-bound pass) per world size for --overlap auto, from the sweeps in profiles/r01_overlap.txt
OVERLAP_AUTO = {2: (4, 62), 4: (4, 62), 8: (4, 62)}
METRIC = "lde_blake2s_commit_throughput"
UNIT = "M31 trace elements/s"
P = 2147483647


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--log-size", type=int, default=LOG_SIZE)
    ap.add_argument("--cols", type=int, default=N_COLS)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--chunks", type=int, default=4, help="column groups per rank of --exchange ce")
    ap.add_argument("--exchange", default="peer", choices=["peer", "ce", "nccl"],
                    help="N > 1: 'peer' = last FFT pass stores into the peers' leaf buffers over NVLink; 'nccl' = all-to-all")
    ap.add_argument("--overlap", default="auto", help="N > 1, --exchange peer: 'groups,peer_sms' = column groups and SMs given to the "
                    "NVLink-bound pass running beside the next group's transforms; 'off' = one group; 'auto' = per world size")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-sample-cols", type=int, default=64)
    ap.add_argument("--fft-radix", type=int, default=0, help="tuning: rows per thread of the FFT passes (8 or 16)")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """Samples SM clocks / throttle reasons during the timed region (B200_PROFILING.md recipe).  NVML is polled from a thread
    every few ms (a multi-GPU timed region lasts tens of ms, shorter than one nvidia-smi -lms period); nvidia-smi is the
    fallback when NVML cannot be loaded."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index=0, uuid=None):
        self.index, self.rows, self.proc, self.nv, self.h, self._stop, self.src = index, [], None, None, None, False, None
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = None
            if uuid:
                try:
                    h = nv.nvmlDeviceGetHandleByUUID(("GPU-" + str(uuid)) if not str(uuid).startswith("GPU-") else str(uuid))
                except Exception:
                    h = None
            if h is None:
                vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
                phys = int(vis.split(",")[index]) if vis and all(x.strip().isdigit() for x in vis.split(",")) else index
                h = nv.nvmlDeviceGetHandleByIndex(phys)
            nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
            self.nv, self.h = nv, h
        except Exception:
            self.nv = None

    def start(self):
        if self.nv:
            self.src = "nvml"
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        try:
            self.src = "nvidia-smi"
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _poll(self):
        nv, h = self.nv, self.h
        masks = [nv.nvmlClocksEventReasonHwSlowdown, nv.nvmlClocksEventReasonHwThermalSlowdown,
                 nv.nvmlClocksEventReasonSwThermalSlowdown, nv.nvmlClocksEventReasonSwPowerCap]
        try:
            mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
        except Exception:
            mx = 0
        while not self._stop:
            try:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                pw = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
                self.rows.append([str(sm), str(mx), "%.2f" % pw] + ["Active" if r & m else "Not Active" for m in masks])
            except Exception:
                pass
            time.sleep(0.003)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def mark(self):
        """Start of the timed region: earlier samples (warm-up) are dropped."""
        self.rows = []

    def stop(self):
        if self.nv:
            self._stop = True
            self.t.join(timeout=1)
        elif not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        else:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                pass
        rows = list(self.rows)
        sm = sorted(int(r[0]) for r in rows if r and r[0].isdigit())
        reasons = set()
        for r in rows:
            for k, nm in enumerate(self.NAMES):
                if len(r) > 3 + k and r[3 + k].lower().startswith("active"):
                    reasons.add(nm)
        mx = [int(r[1]) for r in rows if len(r) > 1 and r[1].isdigit()]
        pw = [float(r[2]) for r in rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        # "under load": drop samples below half of the largest seen (idle ramp)
        load = [s for s in sm if s >= 0.5 * sm[-1]] if sm else []
        return {"sm_mhz": load[len(load) // 2] if load else None, "sm_max_mhz": mx[0] if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons), "source": self.src}


# ------------------------------------------------------------------------------------------------ reference arm
def run_reference(args):
    import numpy as np
    import oracle as orc
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    orc.use_all_cores()
    log, blow, cols = args.log_size, LOG_BLOWUP, args.cpu_sample_cols
    root_init, root_log = orc.canonic_half_coset(log + blow)
    tw, itw = orc.precompute_twiddles(root_init, root_log)
    rng = np.random.default_rng(0x5EED0003)
    trace = rng.integers(0, P, (cols, 1 << log), dtype=np.uint32)
    for _ in range(min(args.warmup, 1)):
        orc.lde_commit(trace, blow, tw, itw)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        orc.lde_commit(trace, blow, tw, itw)
    dt = (time.perf_counter() - t0) / args.steps
    val = cols * (1 << log) / dt
    sample = f"{cols} of {args.cols} columns at log_size {log}, blowup 2 (interpolate + evaluate + Merkle commit of those columns)"
    print(json.dumps({
        "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u32 (M31)",
        "data": "synthetic", "impl": "reference",
        "config": {"workload": f"{args.cols} cols x 2^{log}, blowup 2, Blake2s Merkle commit", "sample": sample},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": orc.num_threads(), "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference = C restatement of TStwo CpuBackend (oracle/); the TypeScript reference needs bun/node, absent here",
    }))


# ------------------------------------------------------------------------------------------------ cuda arm
# Merkle root of the default job (256 x 2^22, blow-up 2, the trace of make_trace): equal at every world size, printed by
# every round's driver run, and checked against the ORACLE's root of the same trace by
# tests/test_gpu_config_sizes.py::test_config3_every_column_and_root_against_oracle
BENCH_ROOT = "6625eaa9c95ec2ba1258dc34cd187601c8eae7333050c25a7a7cb66b3cd3a53d"


def make_trace(torch, dev, rank, world, n_cols, n):
    """This rank's column shard of the synthetic trace: uniform random canonical M31, generated on the device
    in 8 column blocks with fixed seeds so that every world size <= 8 sees the same job (same root)."""
    nblk = 8 if n_cols % 8 == 0 else 1
    bs = n_cols // nblk
    lc = n_cols // world
    assert lc % bs == 0
    trace = torch.empty((lc, n), dtype=torch.int32, device=dev)
    for k in range(lc // bs):
        g = torch.Generator(device=dev); g.manual_seed(0x5EED0003 + rank * (lc // bs) + k)
        trace[k * bs:(k + 1) * bs] = torch.randint(0, P, (bs, n), dtype=torch.int32, device=dev, generator=g)
    return trace


def cpu_baseline(args):
    import numpy as np
    import oracle as orc
    orc.use_all_cores()
    log, blow, cols = args.log_size, LOG_BLOWUP, args.cpu_sample_cols
    root_init, root_log = orc.canonic_half_coset(log + blow)
    tw, itw = orc.precompute_twiddles(root_init, root_log)
    trace = np.random.default_rng(1).integers(0, P, (cols, 1 << log), dtype=np.uint32)
    t0 = time.perf_counter()
    reps = 0
    while reps < 1 or (time.perf_counter() - t0 < 10 and reps < 4):
        orc.lde_commit(trace, blow, tw, itw)
        reps += 1
    dt = (time.perf_counter() - t0) / reps
    return {"value": cols * (1 << log) / dt, "unit": UNIT, "cores": orc.num_threads(), "kind": "port",
            "sample": f"{cols} of {args.cols} columns at log_size {log}, blowup 2, {reps} rep(s), {dt:.2f} s each"}


def run_cuda(args):
    # NCCL prints its version banner on stdout when NCCL_DEBUG=VERSION: the driver wants exactly one JSON line there
    # ("NCCL version ..." appears there on this image): everything written to fd 1 from here on goes to stderr, and
    # rank 0 prints the JSON line on the saved original stdout
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    import numpy as np
    import torch
    import torch.distributed as dist
    from tstwo_b200 import _lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run for --gpus > 1")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    L = _lib.init(local)
    chk = _lib.check
    # a dedicated (non-default) stream shared by torch and the library, so torch.cuda.Event sees the library's kernels
    stream = torch.cuda.Stream(device=local)
    torch.cuda.set_stream(stream)
    chk(L.tstwo_set_stream(C.c_void_p(stream.cuda_stream)))
    if args.fft_radix:
        L.tstwo_debug_set_fft_radix(args.fft_radix)

    log, blow, n_cols = args.log_size, LOG_BLOWUP, args.cols
    ext = log + blow
    n, ne = 1 << log, 1 << ext
    assert n_cols % world == 0 and world & (world - 1) == 0
    lc = n_cols // world                      # local columns
    dev = torch.device("cuda", local)

    tw = C.c_uint64()
    chk(L.tstwo_twiddles_precompute(1 << (31 - (ext - 1) - 2), ext - 1, C.byref(tw)))

    trace = make_trace(torch, dev, rank, world, n_cols, n)
    root = (C.c_uint8 * 32)()
    if world == 1:
        coeffs = torch.empty((lc, n), dtype=torch.int32, device=dev)
        evals = torch.empty((lc, ne), dtype=torch.int32, device=dev)

    if world == 1:
        layers = torch.empty(((2 << ext) - 1) * 8, dtype=torch.int32, device=dev)

        def step():
            chk(L.tstwo_lde_commit(trace.data_ptr(), n_cols, log, blow, tw.value, coeffs.data_ptr(), evals.data_ptr(),
                                   layers.data_ptr(), root, 0))
        launches_per_step = None
    else:
        from tstwo_b200.distributed import CudaShardOps, PeerShardedCommit, ShardedCommit
        if args.exchange in ("peer", "ce"):
            ov = OVERLAP_AUTO.get(world) if args.overlap == "auto" else None if args.overlap == "off" else tuple(int(x) for x in args.overlap.split(","))
            if args.exchange == "ce":
                ov = None
            sc = PeerShardedCommit(local, stream, n_cols, log, blow, overlap=ov)
        else:
            sc = ShardedCommit(CudaShardOps(local, stream), n_cols, log, blow)
        trace_flat = trace.view(-1)
        holder = {}

        def step():
            r = sc.step_pipelined(trace_flat, args.chunks) if args.exchange == "ce" else sc.step(trace_flat)
            if r is not None:
                holder["root"] = r
            else:
                chk(L.tstwo_sync())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing
    try:
        uuid = torch.cuda.get_device_properties(local).uuid
    except Exception:
        uuid = None
    sampler = ClockSampler(local, uuid)
    # one process per GPU: run on the CPUs of the GPU's own NUMA node, so that the pinned host shard of the end-to-end leg is
    # allocated next to the PCIe root port it is read through (8 ranks pulling through one socket's memory halve the rate)
    numa_bound = False
    if world > 1 and sampler.nv is not None and not os.environ.get("TSTWO_NO_NUMA_BIND"):
        try:
            sampler.nv.nvmlDeviceSetCpuAffinity(sampler.h)
            numa_bound = True
        except Exception:
            numa_bound = False
    if rank == 0:
        sampler.start()          # nvidia-smi needs ~100s of ms to start; idle/ramp samples are filtered in stop()
    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    if rank == 0:
        sampler.mark()           # keep only samples taken during the timed region
    l0 = L.tstwo_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1) / args.steps
    launches = int(L.tstwo_launch_count() - l0)
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = n_cols * n / (ms * 1e-3)
    root_hex = bytes(root).hex() if world == 1 else (holder.get("root", b"").hex() if rank == 0 else "")

    # ---- per-kernel timing (1 GPU): the roofline of the dominant kernel, measured live with CUDA events
    roof, kernels, int_roof = None, None, None
    if world == 1:
        kernels = kernel_breakdown(L, chk, torch, stream, trace, coeffs, evals, layers, tw, n_cols, log, blow)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak, which = (peaks.get("hbm_gbs"), "measured (MEASURED_PEAKS.json hbm_gbs)") if peaks.get("hbm_gbs") else (6650.0, "fallback")
        # dominant kernel = the longest single launch (a transform is 2 launches, the leaf layer is 1)
        top_k = max(kernels, key=lambda k: kernels[k]["ms"] / kernels[k]["launches"])
        kk = kernels[top_k]
        ms_launch, by_launch = kk["ms"] / kk["launches"], kk["bytes"] / kk["launches"]
        roof = {"bound": "hbm", "kernel": top_k, "achieved": by_launch / ms_launch / 1e6, "peak": peak, "unit": "GB/s",
                "frac": by_launch / ms_launch / 1e6 / peak, "traffic": NCU_TRAFFIC.get(top_k) if (n_cols, log, blow) == (N_COLS, LOG_SIZE, LOG_BLOWUP) else None, "peak_source": which,
                "algorithmic_bytes_per_launch": by_launch, "ms_per_launch": ms_launch,
                "note": "this kernel is integer-pipe bound, not HBM bound (see int_roofline); traffic = dram bytes read+written "
                        "per launch from profiles/r01_launches_v.txt"}
        # integer-pipe rooflines measured on this pool with tools/microbench*.cu (profiles/r01_microbench_*.txt):
        # 3.667e12 M31 butterflies/s (ALU-pipe bound, 10.15 clk per warp-butterfly per SM sub-partition) and
        # 27.7e9 Blake2s compressions/s (1345 clk per warp-compression per sub-partition, message in registers) at 1965 MHz
        n_bf = n_cols * (log * (n // 2) + log * (ne // 2))          # the top log_blowup layers of evaluate are copies
        n_cmp = ne * ((4 * n_cols + 63) // 64) + ne - 1
        t_fft = kernels["interpolate (2 fft launches)"]["ms"] + kernels["evaluate (2 fft launches)"]["ms"]
        t_mk = kernels["merkle_leaf_layer"]["ms"] + 2 * kernels["merkle_node_layer_top"]["ms"]   # node layers halve: sum ~ 2x the top one
        int_roof = {"fft": {"butterflies": n_bf, "ms": t_fft, "achieved": n_bf / t_fft / 1e6, "peak": 3667.0, "unit": "G butterflies/s",
                            "frac": n_bf / t_fft / 1e6 / 3667.0},
                    "blake2s": {"compressions": n_cmp, "ms": t_mk, "achieved": n_cmp / t_mk / 1e6, "peak": 27.7, "unit": "G compressions/s",
                                "frac": n_cmp / t_mk / 1e6 / 27.7},
                    "peak_source": "tools/microbench.cu, tools/microbench_bf.cu on B200 (profiles/r01_microbench_int_pipes.txt, r01_microbench_stage.txt, r01_microbench_blake2s_forms.txt)"}
        # the dominant kernel's own integer-pipe roofline beside its HBM one (what actually bounds it)
        if top_k == "merkle_leaf_layer":
            n_leaf = ne * ((4 * n_cols + 63) // 64)
            roof["int_pipe"] = {"achieved": n_leaf / ms_launch / 1e6, "peak": 27.7, "unit": "G compressions/s", "frac": n_leaf / ms_launch / 1e6 / 27.7}
        else:
            n_k = n_cols * log * ((ne if top_k.startswith("evaluate") else n) // 2) / kk["launches"]
            roof["int_pipe"] = {"achieved": n_k / ms_launch / 1e6, "peak": 3667.0, "unit": "G butterflies/s", "frac": n_k / ms_launch / 1e6 / 3667.0}

    # ---- end to end through the C ABI with HOST buffers (1 GPU)
    e2e = None
    if world == 1 and not args.no_e2e:
        host = torch.empty((n_cols, n), dtype=torch.int32, pin_memory=True)
        host.copy_(trace.cpu())
        stage = torch.empty((n_cols, n), dtype=torch.int32, device=dev)
        root2 = (C.c_uint8 * 32)()

        def e2e_step():
            chk(L.tstwo_lde_commit_host(host.data_ptr(), n_cols, log, blow, tw.value, stage.data_ptr(), coeffs.data_ptr(),
                                        evals.data_ptr(), layers.data_ptr(), root2, 0))
        for _ in range(2):
            e2e_step()
        torch.cuda.synchronize()
        k = max(3, min(args.steps, 5))
        t0 = time.perf_counter()
        for _ in range(k):
            e2e_step()          # synchronises inside (root is host-visible on return)
        dt = (time.perf_counter() - t0) / k
        assert bytes(root2).hex() == root_hex, "e2e root differs from device-resident root"
        e2e = {"value": n_cols * n / dt, "unit": UNIT, "ms_per_step": dt * 1e3, "h2d_bytes_per_step": n_cols * n * 4,
               "d2h_bytes_per_step": 32, "api": "tstwo_lde_commit_host (C ABI, pinned host trace -> root)"}

    # ---- end to end at N > 1: every rank's column shard starts in pinned HOST memory and crosses its own PCIe link inside the
    # timed region (in column groups, overlapped with the transforms and the NVLink exchange); rank 0 reads the root back
    if world > 1 and not args.no_e2e and args.exchange == "peer":
        host = torch.empty((lc, n), dtype=torch.int32, pin_memory=True)
        host.copy_(trace.cpu())
        stage = torch.empty((lc, n), dtype=torch.int32, device=dev)
        got = {}

        def e2e_step():
            r = sc.step_from_host(host, stage)
            if r is not None:
                got["root"] = r
            else:
                chk(L.tstwo_sync())
        for _ in range(3):
            e2e_step()
        k = max(3, min(args.steps, 5))
        barrier()
        t0 = time.perf_counter()
        for _ in range(k):
            e2e_step()
        barrier()
        dt = torch.tensor([(time.perf_counter() - t0) / k], device=dev, dtype=torch.float64)
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        dt = float(dt.item())
        if rank == 0:
            assert got["root"].hex() == root_hex, "e2e root differs from device-resident root"
        e2e = {"value": n_cols * n / dt, "unit": UNIT, "ms_per_step": dt * 1e3, "h2d_bytes_per_step": n_cols * n * 4,
               "d2h_bytes_per_step": 32,
               "api": f"PeerShardedCommit.step_from_host (pinned host shard per rank -> root on rank 0), {world} PCIe links",
               "cpu_affinity": "NUMA node of each rank's GPU (nvmlDeviceSetCpuAffinity)" if numa_bound else "unbound"}

    if rank == 0:
        cpu = None if args.no_cpu or world > 1 else cpu_baseline(args)
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u32 (M31)",
            "data": "synthetic",
            "config": {"workload": f"{n_cols} cols x 2^{log} M31, blowup 2 (2^{ext} rows), Blake2s Merkle commit",
                       "n_cols": n_cols, "log_size": log, "log_blowup": blow,
                       "l2": "inputs (%.1f GiB) and outputs exceed the 126 MB L2; no flush needed" % (n_cols * n * 4 / 2**30),
                       "parallelism": "1 GPU" if world == 1 else (f"column-sharded LDE, last FFT pass stores rows to their owner over NVLink peer memory -> row-sharded Merkle x{world}"
                                                                       if args.exchange == "peer" else f"column-sharded LDE -> NCCL all-to-all -> row-sharded Merkle x{world}")},
            "gpu_launches": launches, "clocks": clocks, "root": root_hex,
        }
        if roof: out["roofline"] = roof
        if int_roof: out["int_roofline"] = int_roof
        if kernels: out["kernels"] = kernels
        if e2e: out["e2e"] = e2e
        if cpu: out["cpu_baseline"] = cpu
        print(json.dumps(out), file=_REAL_STDOUT, flush=True)
    if world > 1:
        if hasattr(sc, "close"):
            sc.close()
        dist.destroy_process_group()


def kernel_breakdown(L, chk, torch, stream, trace, coeffs, evals, layers, tw, n_cols, log, blow):
    """Times the three phases of the step separately (CUDA events on the launching stream)."""
    from tstwo_b200 import _lib
    ext = log + blow
    n, ne = 1 << log, 1 << ext

    def timed(fn, reps=5):
        fn(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(reps):
            fn()
        b.record(stream); torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    tp = _lib.dptr_array([trace.data_ptr() + 4 * n * c for c in range(n_cols)])
    cp = _lib.dptr_array([coeffs.data_ptr() + 4 * n * c for c in range(n_cols)])
    ep = _lib.dptr_array([evals.data_ptr() + 4 * ne * c for c in range(n_cols)])
    hi_small, hi_big = 1 << (31 - (log - 1) - 2), 1 << (31 - (ext - 1) - 2)
    t_int = timed(lambda: chk(L.tstwo_interpolate_columns(tp, cp, n_cols, log, hi_small, tw.value, 0)))
    t_ev = timed(lambda: chk(L.tstwo_evaluate_columns(cp, log, ep, n_cols, ext, hi_big, tw.value, 0)))
    leaf = layers.data_ptr() + 32 * ((1 << ext) - 1)
    t_leaf = timed(lambda: chk(L.tstwo_merkle_commit_layer(ext, 0, ep, n_cols, leaf)))
    node = layers.data_ptr() + 32 * ((1 << (ext - 1)) - 1)
    t_node = timed(lambda: chk(L.tstwo_merkle_commit_layer(ext - 1, leaf, None, 0, node)))
    b_int, b_ev = 8 * n_cols * n, (4 + 4 * (1 << blow)) * n_cols * n
    b_leaf, b_node = (4 * n_cols + 32) * ne, 96 * (ne // 2)
    mk = lambda ms, by, launches: {"ms": ms, "bytes": by, "gbs": by / ms / 1e6, "launches": launches}
    return {"interpolate (2 fft launches)": mk(t_int, b_int, 2), "evaluate (2 fft launches)": mk(t_ev, b_ev, 2),
            "merkle_leaf_layer": mk(t_leaf, b_leaf, 1), "merkle_node_layer_top": mk(t_node, b_node, 1)}


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_cuda(a)

#!/usr/bin/env python
"""bench.py -- LDE + Blake2s-commit throughput (BASELINE.json metric) on N B200s of one node.

A step = one pass of the hot path over one batch of synthetic input: interpolate 256 trace columns of 2^22 M31
elements, evaluate them on the blow-up-2 canonic domain (2^23 rows), Blake2s Merkle commit, root on the host.
  value : trace elements / s with the trace resident in HBM (CUDA events, max over ranks)
  e2e   : the same through the C-ABI call a host makes (tstwo_lde_commit_host): trace in pinned HOST memory,
          host->device copies inside the timed region, root read back
N > 1 : columns are sharded by column for the LDE; the last FFT pass stores every row range into its owner's leaf buffer over
        NVLink peer memory (--exchange nccl: an all-to-all instead), the ranks order the step with release/acquire flags in peer
        memory (--fence nccl: collectives), every rank hashes its leaves and builds its subtree, rank 0 finishes the top log2(N)
        layers.  --ownership halves (default): a rank holds its share of the first half of the columns and of the second half, so
        that the first half of every leaf is hashed while the exchange still runs (contiguous: one column range per rank, leaves
        hashed after the exchange); the job and its root are the same.  --single-process drives the N devices from ONE host thread through tstwo_lde_commit_multi_dev.
--impl reference : the CPU oracle (C restatement of TStwo's CpuBackend; the TypeScript reference cannot run here)
        on all host cores, on the whole workload (all 256 columns; --cpu-sample-cols bounds it).
roofline / rooflines / int_roofline / e2e.h2d_only / extras / phases: see DESIGN.md section 6.
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
# ncu evidence the roofline entries cite: per-kernel means of the launch list of THIS command at the headline size
# (scripts/launch_list.py --json) and the --set full counters of the FFT passes (scripts/ncu_summary.py --json)
NCU_LAUNCHES = "profiles/r02_launches.json"
NCU_COUNTERS = "profiles/r02_fft_ncu.json"
# (column groups, SMs of the NVLink-bound pass) per world size for --overlap auto, from the sweeps in profiles/r01_overlap.txt
OVERLAP_AUTO = {2: (4, 62), 4: (4, 62), 8: (4, 66)}    # profiles/r01_overlap.txt, r02_multi.txt
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
    ap.add_argument("--fence", default="flags", choices=["flags", "nccl"],
                    help="N > 1, --exchange peer: cross-rank ordering by release/acquire flags in peer memory (default) or NCCL collectives")
    ap.add_argument("--ownership", default="halves", choices=["contiguous", "halves"],
                    help="N > 1, --exchange peer with flags: which trace columns a rank holds (tstwo_b200.distributed.owned_columns); "
                         "'halves' lets the receivers hash the first half of every leaf beside the last storing pass")
    ap.add_argument("--e2e-chunks", type=int, default=8, help="N > 1 end-to-end leg: column groups per rank crossing PCIe (the last group's LDE and the Merkle tree are exposed)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-sample-cols", type=int, default=0, help="columns of the CPU legs (0 = all of --cols: the whole config)")
    ap.add_argument("--no-extras", action="store_true", help="skip the config 4 (FRI) and config 5 (prove) timings")
    ap.add_argument("--single-process", action="store_true", help="N GPUs driven by ONE process through tstwo_lde_commit_multi")
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
def _cpu_leg_setup(args):
    """The CPU legs (reference arm, cpu_baseline): the oracle rebuilt -O3 -march=native ON THIS HOST, all its cores, the WHOLE
    config (every column) unless --cpu-sample-cols asks for a sample."""
    import oracle as orc
    native = orc.use_native_build()
    orc.use_all_cores()
    cols = args.cpu_sample_cols if args.cpu_sample_cols > 0 else args.cols
    return orc, min(cols, args.cols), native


def run_reference(args):
    import numpy as np
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    orc, cols, native = _cpu_leg_setup(args)
    log, blow = args.log_size, LOG_BLOWUP
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
    whole = cols == args.cols
    sample = (f"all {cols} columns" if whole else f"{cols} of {args.cols} columns") + \
        f" at log_size {log}, blowup 2 (interpolate + evaluate + Merkle commit), oracle built {'-O3 -march=native on this host' if native else '-O3 -march=x86-64-v2'}"
    print(json.dumps({
        "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u32 (M31)",
        "data": "synthetic", "impl": "reference",
        "config": {"workload": f"{args.cols} cols x 2^{log} M31, blowup 2 (2^{log + blow} rows), Blake2s Merkle commit",
                   "n_cols": args.cols, "log_size": log, "log_blowup": blow, "same_config": whole, "sample": sample},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": orc.num_threads(), "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference = C restatement of TStwo CpuBackend (oracle/); the TypeScript reference needs bun/node, absent here",
    }))


# ------------------------------------------------------------------------------------------------ cuda arm
# Merkle root of the default job (256 x 2^22, blow-up 2, the trace of make_trace): equal at every world size, printed by
# every round's driver run, and checked against the ORACLE's root of the same trace by
# tests/test_gpu_config_sizes.py::test_config3_every_column_and_root_against_oracle
BENCH_ROOT = "6625eaa9c95ec2ba1258dc34cd187601c8eae7333050c25a7a7cb66b3cd3a53d"


def make_trace(torch, dev, rank, world, n_cols, n, ownership="contiguous"):
    """This rank's column shard of the synthetic trace: uniform random canonical M31, generated on the device
    in 8 column blocks with fixed seeds (block b of the TRACE has seed 0x5EED0003 + b) so that every world size <= 8 and either
    column ownership (tstwo_b200.distributed.owned_columns) sees the same job (same root)."""
    from tstwo_b200.distributed import owned_columns
    nblk = 8 if n_cols % 8 == 0 else 1
    bs = n_cols // nblk
    lc = n_cols // world
    trace = torch.empty((lc, n), dtype=torch.int32, device=dev)
    at = 0
    for first, count in owned_columns(rank, world, n_cols, ownership):
        c = first
        while c < first + count:
            b, off = divmod(c, bs)
            take = min(bs - off, first + count - c)
            g = torch.Generator(device=dev); g.manual_seed(0x5EED0003 + b)
            blk = torch.randint(0, P, (bs, n), dtype=torch.int32, device=dev, generator=g)
            trace[at:at + take] = blk[off:off + take]
            del blk
            at += take
            c += take
    return trace


def cpu_baseline(args):
    import numpy as np
    orc, cols, native = _cpu_leg_setup(args)
    log, blow = args.log_size, LOG_BLOWUP
    root_init, root_log = orc.canonic_half_coset(log + blow)
    tw, itw = orc.precompute_twiddles(root_init, root_log)
    trace = np.random.default_rng(1).integers(0, P, (cols, 1 << log), dtype=np.uint32)
    t0 = time.perf_counter()
    reps = 0
    while reps < 1 or (time.perf_counter() - t0 < 10 and reps < 3):
        orc.lde_commit(trace, blow, tw, itw)
        reps += 1
    dt = (time.perf_counter() - t0) / reps
    return {"value": cols * (1 << log) / dt, "unit": UNIT, "cores": orc.num_threads(), "kind": "port",
            "sample": (f"all {cols} columns" if cols == args.cols else f"{cols} of {args.cols} columns") +
                      f" at log_size {log}, blowup 2, {reps} rep(s), {dt:.2f} s each, oracle built "
                      f"{'-O3 -march=native on this host' if native else '-O3 -march=x86-64-v2'}"}


def run_single_process(args):
    """N GPUs driven by ONE process / one host thread through tstwo_init_devices + tstwo_lde_commit_multi[_dev] (the form a
    TStwo host can call: SURVEY 8b).  Same job, same trace, same JSON keys as the torchrun arm."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    import torch
    from tstwo_b200 import _lib
    G = args.gpus
    L = _lib.load(); chk = _lib.check
    chk(L.tstwo_init_devices(G, (C.c_int * G)(*range(G))))
    log, blow, n_cols = args.log_size, LOG_BLOWUP, args.cols
    n = 1 << log
    lc = n_cols // G
    shards = [make_trace(torch, torch.device("cuda", d), d, G, n_cols, n) for d in range(G)]
    for d in range(G):
        torch.cuda.synchronize(d)
    ptrs = (C.c_uint64 * G)(*[s.data_ptr() for s in shards])
    root = (C.c_uint8 * 32)()
    step = lambda: chk(L.tstwo_lde_commit_multi_dev(ptrs, n_cols, log, blow, root, 0))
    sampler = ClockSampler(0, None)
    sampler.start()
    for _ in range(max(args.warmup, 3)):
        step()
    sampler.mark()
    chk(L.tstwo_set_device(0))
    l0 = L.tstwo_launch_count()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()                       # returns with the root on the host: every device has finished
    ms = (time.perf_counter() - t0) / args.steps * 1e3
    launches = 0
    for d in range(G):
        chk(L.tstwo_set_device(d)); launches += int(L.tstwo_launch_count())
    chk(L.tstwo_set_device(0))
    clocks = sampler.stop()
    root_hex = bytes(root).hex()
    e2e = None
    if not args.no_e2e:
        host = torch.empty((n_cols, n), dtype=torch.int32, pin_memory=True)
        for d in range(G):
            host[d * lc:(d + 1) * lc].copy_(shards[d].cpu())
        del shards
        root2 = (C.c_uint8 * 32)()
        hstep = lambda: chk(L.tstwo_lde_commit_multi(C.c_void_p(host.data_ptr()), n_cols, log, blow, root2, 0))
        for _ in range(3):
            hstep()
        k = max(3, min(args.steps, 5))
        t0 = time.perf_counter()
        for _ in range(k):
            hstep()
        dt = (time.perf_counter() - t0) / k
        assert bytes(root2).hex() == root_hex, "e2e root differs from device-resident root"
        e2e = {"value": n_cols * n / dt, "unit": UNIT, "ms_per_step": dt * 1e3, "h2d_bytes_per_step": n_cols * n * 4, "d2h_bytes_per_step": 32,
               "api": f"tstwo_lde_commit_multi (C ABI, one host thread, pinned host trace -> root), {G} PCIe links"}
    out = {"metric": METRIC, "value": n_cols * n / (ms * 1e-3), "unit": UNIT, "n_gpus": G, "steps": args.steps, "warmup": max(args.warmup, 3),
           "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u32 (M31)", "data": "synthetic",
           "config": {"workload": f"{n_cols} cols x 2^{log} M31, blowup 2 (2^{log + blow} rows), Blake2s Merkle commit", "n_cols": n_cols,
                      "log_size": log, "log_blowup": blow,
                      "parallelism": f"ONE process, one host thread, {G} devices: tstwo_init_devices + tstwo_lde_commit_multi_dev "
                                     "(column-sharded LDE, last FFT pass stores rows to their owner over NVLink peer access, events as fences)",
                      "timing": "host wall clock around calls that return with the root on the host (every device done); "
                                "launches = sum over devices"},
           "gpu_launches": launches - int(l0), "clocks": clocks, "root": root_hex,
           "root_matches_oracle_checked_root": root_hex == BENCH_ROOT if (n_cols, log, blow) == (N_COLS, LOG_SIZE, LOG_BLOWUP) else None}
    if e2e: out["e2e"] = e2e
    print(json.dumps(out), file=_REAL_STDOUT, flush=True)
    chk(L.tstwo_shutdown())


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

    ownership = "contiguous"
    if world > 1 and args.exchange == "peer" and args.ownership == "halves":
        from tstwo_b200.distributed import PeerShardedCommit as _PSC
        _ov = OVERLAP_AUTO.get(world) if args.overlap == "auto" else None if args.overlap == "off" else tuple(int(x) for x in args.overlap.split(","))
        if _PSC.halves_supported(world, n_cols, log, _ov, args.fence):
            ownership = "halves"
    trace = make_trace(torch, dev, rank, world, n_cols, n, ownership)
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
            sc = PeerShardedCommit(local, stream, n_cols, log, blow, overlap=ov, fence=args.fence,
                                   ownership=ownership if args.exchange == "peer" else "contiguous")
        else:
            sc = ShardedCommit(CudaShardOps(local, stream), n_cols, log, blow)
        trace_flat = trace.view(-1)
        holder = {}

        # device-resident steps are enqueued back to back (no host synchronisation inside the timed region: the ranks order
        # themselves with flags in peer memory); the root is read once after the loop
        def step():
            if args.exchange == "peer":
                sc.step(trace_flat, read_root=False)
                return
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
    if world > 1 and args.exchange == "peer" and rank == 0:
        holder["root"] = sc.root_bytes()
    root_hex = bytes(root).hex() if world == 1 else (holder.get("root", b"").hex() if rank == 0 else "")

    # ---- per-rank phase breakdown (N > 1, fused peer stores): CUDA events at the phase boundaries, min / max over ranks, and
    # the NVLink rate of the pass that stores the evaluations into the peers' leaf buffers
    phases = None
    if world > 1 and args.exchange == "peer":
        recs = []
        for _ in range(2):
            sc.step_phases(trace_flat)
        barrier()
        for _ in range(6):
            recs.append(sc.step_phases(trace_flat))
        torch.cuda.synchronize()
        names = ["lde_and_peer_stores", "fence", "leaf_layer_and_subtree", "root_gather_and_top_layers"]
        if ownership == "halves":                          # same boundaries; what lies between them differs
            names = ["lde_and_peer_stores_and_first_half_of_leaves", "fence", "second_half_of_leaves_and_subtree", "root_gather_and_top_layers"]
        tot = [sum(r[i].elapsed_time(r[i + 1]) for r in recs) / len(recs) for i in range(4)]
        t = torch.tensor(tot, device=dev, dtype=torch.float64)
        mx, mn = t.clone(), t.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX); dist.all_reduce(mn, op=dist.ReduceOp.MIN)
        sent = lc * ne * 4 * (world - 1) / world           # bytes this rank stores into peer memory per step
        phases = {nm: {"ms_min_over_ranks": float(mn[i]), "ms_max_over_ranks": float(mx[i])} for i, nm in enumerate(names)}
        phases["nvlink"] = {"bytes_stored_to_peers_per_rank": sent, "gbs_per_rank_over_the_lde_phase": sent / (float(mx[0]) * 1e-3) / 1e9,
                            "reference": "770 GB/s per direction measured peer copy (B200_PROFILING.md), 900 nominal"}
        phases["note"] = ("CUDA events on each rank's stream; a rank's fence / gather time includes waiting for the slowest peer. "
                          "single-GPU pure LDE of the same shard = kernels.passes of the N=1 line / N")
        # the same roofline entries as the N = 1 line, per rank: the LDE phase (the four FFT passes of the rank's column shard, the
        # last one storing into the peers) and the leaf layer + subtree of the rank's row shard, against the measured HBM peak
        try:
            pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs") or 6650.0
        except Exception:
            pk = 6650.0
        el = lc * n
        by_lde, by_merkle = (8 + 4 + 4 * (1 << blow)) * el, (4 * n_cols + 32 + 96) * (ne // world)
        mk = lambda name, by, t_ms, note: {"kernel": name, "bound": "hbm", "algorithmic_bytes": by, "ms": t_ms, "achieved": by / t_ms / 1e6, "peak": pk,
                                           "unit": "GB/s", "frac": by / t_ms / 1e6 / pk, "traffic": None, "share_of_step": t_ms / ms, "note": note}
        if ownership == "halves":
            # the first phase also hashes the first half of every leaf (beside the last storing pass): its bytes move with it
            by_half = (4 * (n_cols // 2)) * (ne // world)
            phases["rooflines_per_rank"] = [
                mk("circle FFT of the rank's column shard (4 x column groups x fft_tma_kernel, last pass stores to the peers) + the first half "
                   "of every leaf (merkle_leaf_chunk_kernel beside the last storing pass)", by_lde + by_half, float(mx[0]),
                   "integer-pipe bound as at N = 1; at N = 8 the peer stores (573 GB/s all-to-all) bound this phase (profiles/r02_multi.txt)"),
                mk("second half of every leaf (merkle_leaf_chunk_kernel) + subtree of the rank's row shard", by_merkle - by_half, float(mx[2]),
                   "integer-pipe bound (BLAKE2s)")]
        else:
            phases["rooflines_per_rank"] = [
                mk("circle FFT of the rank's column shard (4 x column groups x fft_tma_kernel, last pass stores to the peers)", by_lde, float(mx[0]),
                   "integer-pipe bound as at N = 1; at N = 8 the peer stores (573 GB/s all-to-all) bound this phase (profiles/r02_multi.txt)"),
                mk("merkle leaf layer + subtree of the rank's row shard", by_merkle, float(mx[2]), "integer-pipe bound (BLAKE2s)")]
        barrier()

    # ---- per-kernel timing (1 GPU), live CUDA events: `roofline` = the phase with the largest share of the step (the four
    # circle-FFT passes), `rooflines` = every kernel of the step, each with its algorithmic bytes, its live
    # duration, the fraction of the measured HBM peak and the DRAM traffic ncu saw for the same command (committed JSON)
    roof, rooflines, kernels, int_roof = None, None, None, None
    if world == 1:
        kernels = kernel_breakdown(L, chk, torch, stream, trace, coeffs, evals, layers, tw, n_cols, log, blow)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak, which = (peaks.get("hbm_gbs"), "measured (MEASURED_PEAKS.json hbm_gbs)") if peaks.get("hbm_gbs") else (6650.0, "fallback (B200_PROFILING.md)")
        headline = (n_cols, log, blow) == (N_COLS, LOG_SIZE, LOG_BLOWUP)
        ncu_l, ncu_c = {}, {}
        if headline:
            try:
                ncu_l = json.load(open(os.path.join(ROOT, NCU_LAUNCHES)))["kernels"]
            except Exception:
                ncu_l = {}
            try:
                ncu_c = json.load(open(os.path.join(ROOT, NCU_COUNTERS)))["kernels"]
            except Exception:
                ncu_c = {}

        def entry(name, ncu_names, ms, by, launches, note=None, mult=None):
            tr = [ncu_l.get(k) for k in ncu_names]
            mult = mult or [1] * len(ncu_names)            # launches of each kernel inside one step (the JSON holds per-launch means)
            traffic = sum(m * (t["dram_read_bytes"] + t["dram_write_bytes"]) for m, t in zip(mult, tr)) if tr and all(tr) else None
            e = {"kernel": name, "bound": "hbm", "launches_per_step": launches, "algorithmic_bytes": by, "ms": ms,
                 "achieved": by / ms / 1e6, "peak": peak, "unit": "GB/s", "frac": by / ms / 1e6 / peak,
                 "traffic": traffic, "traffic_source": NCU_LAUNCHES if traffic is not None else None, "share_of_step": ms / ms_step}
            cs = [ncu_c.get(k) for k in ncu_names]
            if cs and all(cs):
                e["ncu"] = {"alu_pipe_active_pct": [c.get("alu_pipe_active_pct") for c in cs], "issue_active_pct": [c.get("issue_active_pct") for c in cs],
                            "tensor_pipe_active_pct": [c.get("tensor_pipe_active_pct") for c in cs], "source": NCU_COUNTERS}
            if note:
                e["note"] = note
            return e

        ms_step = ms
        el = n_cols * n
        pk = kernels["passes"]
        rooflines = []
        pass_bytes = {"inverse contiguous": 8 * el, "inverse strided": 8 * el, "forward strided": (4 + 4 * (1 << blow)) * el,
                      "forward contiguous": 8 * (1 << blow) * el}
        for nm, v in pk.items():
            rooflines.append(entry(f"fft pass: {nm} ({v['kernel']})", [v["kernel"]], v["ms"], pass_bytes.get(nm, 0), 1,
                                   "bytes = this pass's own input read once + output written once"))
        t_int, t_ev = kernels["interpolate (2 fft launches)"]["ms"], kernels["evaluate (2 fft launches)"]["ms"]
        names_int = [v["kernel"] for nm, v in pk.items() if nm.startswith("inverse")]
        names_ev = [v["kernel"] for nm, v in pk.items() if nm.startswith("forward")]
        rooflines.append(entry("interpolate (2 fft passes)", names_int, t_int, 8 * el, 2, "SURVEY 8(d): 8 B per trace element"))
        rooflines.append(entry("evaluate, blow-up 2 (2 fft passes)", names_ev, t_ev, (4 + 4 * (1 << blow)) * el, 2, "SURVEY 8(d): 12 B per trace element"))
        t_leaf, t_nodes = kernels["merkle_leaf_layer"]["ms"], kernels["merkle_node_layers"]["ms"]
        rooflines.append(entry("merkle leaf layer (merkle_leaf_tma_kernel)", ["merkle_leaf_tma_kernel"], t_leaf, (4 * n_cols + 32) * ne, 1,
                               "integer-pipe bound (int_roofline.blake2s), not HBM bound"))
        n_node_launches = max(0, ext - 10)                 # layers ext-1 .. 10 one launch each; layers 9 .. 0 in merkle_tail_kernel
        rooflines.append(entry(f"merkle node layers (in a step: {n_node_launches} launches of merkle_node_kernel + merkle_tail_kernel; timed here layer by layer)",
                               ["merkle_node_kernel", "merkle_tail_kernel"], t_nodes, 96 * (ne - 1), n_node_launches + 1,
                               "integer-pipe bound like the leaf layer", mult=[n_node_launches, 1]))
        roof = entry("circle FFT: interpolate + evaluate (4 launches of fft_tma_kernel)", names_int + names_ev, t_int + t_ev,
                     (8 + 4 + 4 * (1 << blow)) * el, 4,
                     "largest share of the step; HBM is the bound SURVEY 8(d) assigns, the passes are integer-pipe / issue limited "
                     "(ncu counters per pass in rooflines[].ncu; int_roofline.fft)")
        # integer-pipe rooflines measured on this pool with tools/microbench*.cu (profiles/r01_microbench_*.txt):
        # 3.667e12 M31 butterflies/s for the 7-instruction butterfly (10.15 clk per warp-butterfly per SM sub-partition), the
        # hardware bound of its 4 ALU-pipe instructions is 8 clk -> 4.65e12/s; 27.7e9 Blake2s compressions/s (1345 clk per
        # warp-compression per sub-partition, message in registers) at 1965 MHz
        n_bf = n_cols * (log * (n // 2) + log * (ne // 2))          # the top log_blowup layers of evaluate are copies
        n_cmp = ne * ((4 * n_cols + 63) // 64) + ne - 1
        t_fft, t_mk = t_int + t_ev, t_leaf + t_nodes
        int_roof = {"fft": {"butterflies": n_bf, "ms": t_fft, "achieved": n_bf / t_fft / 1e6, "peak": 3667.0, "unit": "G butterflies/s",
                            "frac": n_bf / t_fft / 1e6 / 3667.0, "alu_pipe_bound": 4650.0, "frac_of_alu_pipe_bound": n_bf / t_fft / 1e6 / 4650.0,
                            "note": "peak = this library's 7-instruction butterfly in a register-resident microbenchmark; alu_pipe_bound = "
                                    "2 clk x 4 ALU-pipe instructions per warp-butterfly.  8 of the 44 layers run on the tensor cores "
                                    "(tcgen05 kind::i8) and are counted as butterflies all the same."},
                    "blake2s": {"compressions": n_cmp, "ms": t_mk, "achieved": n_cmp / t_mk / 1e6, "peak": 27.7, "unit": "G compressions/s",
                                "frac": n_cmp / t_mk / 1e6 / 27.7},
                    "peak_source": "tools/microbench.cu, tools/microbench_bf.cu on B200 (profiles/r01_microbench_int_pipes.txt, r01_microbench_stage.txt, r01_microbench_blake2s_forms.txt)"}

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
        # the PCIe floor of that step: the same 4 GiB copied from the same pinned buffer with no kernel running
        stage.copy_(host, non_blocking=True); torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(k):
            stage.copy_(host, non_blocking=True)
        torch.cuda.synchronize()
        dh = (time.perf_counter() - t0) / k
        e2e = {"value": n_cols * n / dt, "unit": UNIT, "ms_per_step": dt * 1e3, "h2d_bytes_per_step": n_cols * n * 4,
               "d2h_bytes_per_step": 32, "api": "tstwo_lde_commit_host (C ABI, pinned host trace -> root)",
               "h2d_only": {"ms_per_step": dh * 1e3, "gbs_per_link": n_cols * n * 4 / dh / 1e9,
                            "note": "the same pinned buffer copied to the device with no kernel running: the PCIe floor of the e2e step"},
               "e2e_over_floor": dt / max(dh, ms * 1e-3)}

    # ---- end to end at N > 1: every rank's column shard starts in pinned HOST memory and crosses its own PCIe link inside the
    # timed region (in column groups, overlapped with the transforms and the NVLink exchange); rank 0 reads the root back
    if world > 1 and not args.no_e2e and args.exchange == "peer":
        host = torch.empty((lc, n), dtype=torch.int32, pin_memory=True)
        host.copy_(trace.cpu())
        stage = torch.empty((lc, n), dtype=torch.int32, device=dev)
        got = {}

        def e2e_step():
            r = sc.step_from_host(host, stage, args.e2e_chunks)
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
        # the platform's share of that: ALL ranks copying their pinned shard to their GPU at the same time, no kernel at all
        # (what the PCIe links + host memory deliver to N GPUs at once); e2e cannot beat max(this, the device-resident step)
        def h2d_only():
            stage.copy_(host, non_blocking=True)
        h2d_only()
        barrier()
        t0 = time.perf_counter()
        for _ in range(k):
            h2d_only()
        barrier()
        dh = torch.tensor([(time.perf_counter() - t0) / k], device=dev, dtype=torch.float64)
        dist.all_reduce(dh, op=dist.ReduceOp.MAX)
        dh = float(dh.item())
        e2e = {"value": n_cols * n / dt, "unit": UNIT, "ms_per_step": dt * 1e3, "h2d_bytes_per_step": n_cols * n * 4,
               "d2h_bytes_per_step": 32,
               "api": f"PeerShardedCommit.step_from_host (pinned host shard per rank -> root on rank 0), {world} PCIe links",
               "cpu_affinity": "NUMA node of each rank's GPU (nvmlDeviceSetCpuAffinity)" if numa_bound else "unbound",
               "column_groups_per_rank": args.e2e_chunks,
               "h2d_only": {"ms_per_step": dh * 1e3, "gbs_per_link": lc * n * 4 / dh / 1e9, "gbs_aggregate": n_cols * n * 4 / dh / 1e9,
                            "note": "all ranks copy their pinned shard concurrently, no kernels: the PCIe / host-memory floor of the e2e step"},
               "e2e_over_floor": dt / max(dh, ms * 1e-3)}

    extras = None
    if world == 1 and not args.no_extras:
        del coeffs, evals, layers, trace
        torch.cuda.empty_cache()
        L.tstwo_mem_trim()
        extras = {}
        for key, fn in (("config4_fri", lambda: config4_fri(L, torch)), ("config5_prove", lambda: config5_prove(L, torch))):
            try:
                extras[key] = fn()
            except Exception as e:   # an extra must never cost the headline line
                extras[key] = {"error": f"{type(e).__name__}: {e}"[:300]}
            L.tstwo_mem_trim()
    try:
        sass_patch = open(_lib.SO_PATH + ".sasspatch").read().strip()
    except OSError:
        sass_patch = "unknown"
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
            **({"fence": ("release/acquire flags in peer memory (tstwo_peer_signal / _wait / _put), no collective and no host synchronisation inside a step"
                          if args.fence == "flags" else "NCCL all-reduce + all-gather"),
                "column_ownership": ("halves: a rank holds its share of the trace's first half and of its second half; the receivers hash the first "
                                     "half of every leaf beside the last storing pass (tstwo_merkle_leaf_absorb)" if ownership == "halves"
                                     else "contiguous column ranges; leaves hashed once all columns have arrived")}
               if world > 1 and args.exchange == "peer" else {}),
            "gpu_launches": launches, "clocks": clocks, "root": root_hex, "root_matches_oracle_checked_root": root_hex == BENCH_ROOT if (n_cols, log, blow) == (N_COLS, LOG_SIZE, LOG_BLOWUP) else None,
            "sass_patch": sass_patch,
        }
        if extras: out["extras"] = extras
        if phases:
            out["phases"] = phases
            out["roofline"] = dict(phases["rooflines_per_rank"][0], per="rank")
        if roof: out["roofline"] = roof
        if rooflines: out["rooflines"] = rooflines
        if int_roof: out["int_roofline"] = int_roof
        if kernels: out["kernels"] = kernels
        if e2e: out["e2e"] = e2e
        if cpu: out["cpu_baseline"] = cpu
        print(json.dumps(out), file=_REAL_STDOUT, flush=True)
    if world > 1:
        if hasattr(sc, "close"):
            sc.close()
        dist.destroy_process_group()


def config4_fri(L, torch):
    """BASELINE config 4 as extra keys: the 23 folds of one QM31 column from 2^24 down to 2 elements (kernel time, CUDA events)
    and FriProver.commit (folds + one Merkle tree per layer + transcript; wall clock with a device sync on both sides)."""
    import numpy as np
    import tstwo_b200 as tb
    from tstwo_b200.channel import Blake2sChannel
    from tstwo_b200.fri import FriProver
    log = 24
    rng = np.random.default_rng(0x5EED0004)
    dom = tb.CanonicCoset.new(log).circleDomain()
    tw = tb.CudaBackend.precomputeTwiddles(dom.halfCoset)
    src = tb.SecureEvaluation(dom, tb.SecureColumnByCoords.from_cpu([rng.integers(0, P, 1 << log, dtype=np.uint32) for _ in range(4)]))
    alpha = [int(x) for x in rng.integers(1, P, 4)]

    def chain():
        layer = tb.LineEvaluation.newZero(tb.LineDomain.new(dom.halfCoset))
        tb.fold_circle_into_line(layer, src, alpha, tw)
        while layer.len() > 2:
            layer = tb.fold_line(layer, alpha, tw)
    chain(); L.tstwo_sync()
    best = 1e9
    for _ in range(5):
        L.tstwo_sync(); t0 = time.perf_counter(); chain(); L.tstwo_sync()
        best = min(best, time.perf_counter() - t0)
    n = 1 << log
    by = 16 * n + 32 * (n // 2) + sum(16 * m + 16 * (m // 2) + 4 * (m // 2) for m in (1 << k for k in range(log - 1, 1, -1)))
    coords = [tb.CudaCirclePoly.new(rng.integers(0, P, 1 << (log - 1), dtype=np.uint32)).evaluateWithTwiddles(dom, tw).values for _ in range(4)]
    # a SecureColumnByCoords as the prover's own steps produce it (accumulate_quotients, fold outputs): ONE block, four quarters
    sc = tb.SecureColumnByCoords.uninitialized(1 << log)
    for k in range(4):
        L.tstwo_copy(sc.columns[k].ptr, coords[k].ptr, 4 << log)
    del coords
    col = tb.SecureEvaluation(dom, sc)
    cfg = tb.FriConfig(0, 1, 30)
    FriProver.commit(Blake2sChannel.create(), cfg, [col], tw); L.tstwo_sync()
    t_commit, t_dec = 1e9, 1e9
    for _ in range(3):
        t0 = time.perf_counter(); pr = FriProver.commit(Blake2sChannel.create(), cfg, [col], tw); L.tstwo_sync()
        t_commit = min(t_commit, time.perf_counter() - t0)
        t0 = time.perf_counter(); pr.decommit(Blake2sChannel.create()); L.tstwo_sync()
        t_dec = min(t_dec, time.perf_counter() - t0)
    return {"workload": "one QM31 column on CanonicCoset(24): fold_circle_into_line + 22 fold_line (to 2 elements)",
            "fold_chain_ms": best * 1e3, "fold_chain_algorithmic_bytes": by, "fold_chain_gbs": by / best / 1e9,
            "qm31_elements_per_s": n / best, "fri_commit_ms": t_commit * 1e3, "fri_decommit_30_queries_ms": t_dec * 1e3,
            "fri_commit_compressions": 4 * n, "fri_commit_gcompress_per_s": 4 * n / t_commit / 1e9, "fri_inner_layers": len(pr.inner_layers),
            "timing": "wall clock, device sync on both sides (23 launches; includes launch latency)"}


def config5_prove(L, torch):
    """BASELINE config 5 as extra keys: wide-Fibonacci prove() at log_size 23 with 100 columns (Stwo's FIB_SEQUENCE_LENGTH),
    blow-up 2, 70 queries, 10 proof-of-work bits; second pass (allocator warm).  The trace is built on the device."""
    import tstwo_b200 as tb
    from tstwo_b200 import air, prover
    from tstwo_b200.channel import Blake2sChannel
    log, n_cols, blow = 23, 100, 1
    clock = lambda: (L.tstwo_sync(), time.perf_counter())[1]
    # generateTestTrace (examples/fibonacci.ts:245-258) on the device: row r starts (1, r), then c = a^2 + b^2
    n = 1 << log
    trace = torch.empty((n_cols, n), dtype=torch.int32, device="cuda")
    a = torch.ones(n, dtype=torch.int64, device="cuda")
    b = torch.arange(n, dtype=torch.int64, device="cuda")
    trace[0], trace[1] = a.to(torch.int32), b.to(torch.int32)
    for k in range(2, n_cols):
        a, b = b, (a * a % P + b * b % P) % P
        trace[k] = b.to(torch.int32)
    del a, b
    torch.cuda.synchronize()
    tw = tb.CudaBackend.precomputeTwiddles(tb.CanonicCoset.new(log + 1 + blow).half_coset())
    dom = tb.CanonicCoset.new(log).circleDomain()
    cols = [tb.CudaCircleEvaluation(dom, tb.CudaColumn.view(trace.data_ptr() + 4 * (c << log), n, trace)) for c in range(n_cols)]
    res = None
    for rep in range(4):
        channel = Blake2sChannel.create()
        scheme = tb.CommitmentSchemeProver(tb.PcsConfig(10, tb.FriConfig(0, blow, 70)), tw)
        t3 = clock()
        b = scheme.tree_builder(); b.extend_evals([]); b.commit(channel)
        b = scheme.tree_builder(); b.extend_evals(cols); b.commit(channel)
        t4 = clock()
        comp = air.WideFibonacciComponent(air.WideFibonacciEval.create(log, n_cols))
        proof = prover.prove([comp], channel, scheme)
        t5 = clock()
        if rep and (res is None or t5 - t3 < res["total_ms"] * 1e-3):
            res = {"workload": f"wide-Fibonacci prove(), {n_cols} columns x 2^{log}, blow-up 2, 70 queries, 10 pow bits",
                   "trace_commit_ms": (t4 - t3) * 1e3, "prove_ms": (t5 - t4) * 1e3, "total_ms": (t5 - t3) * 1e3,
                   "trace_elements_per_s": n_cols * (1 << log) / (t5 - t3),
                   "timing": "wall clock, device sync on both sides; best of 3 passes after one warm-up pass (prove() is host-latency bound)"}
        del scheme, proof
    return res


def kernel_breakdown(L, chk, torch, stream, trace, coeffs, evals, layers, tw, n_cols, log, blow):
    """Times the phases of the step separately (CUDA events on the launching stream) and every FFT pass launch on its own
    (tstwo_debug_pass_timing: events recorded by the library around each launch)."""
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
    # every pass launch on its own
    passes = {}
    reps = 5
    L.tstwo_debug_pass_timing(1)
    for _ in range(reps):
        chk(L.tstwo_interpolate_columns(tp, cp, n_cols, log, hi_small, tw.value, 0))
        chk(L.tstwo_evaluate_columns(cp, log, ep, n_cols, ext, hi_big, tw.value, 0))
    tags, mss = (C.c_int * 64)(), (C.c_float * 64)()
    cnt = L.tstwo_debug_pass_times(64, tags, mss)
    L.tstwo_debug_pass_timing(0)
    acc = {}
    for i in range(max(cnt, 0)):
        acc.setdefault(tags[i], []).append(mss[i])
    for tag, v in acc.items():
        nm = ("inverse " if tag & 1 else "forward ") + ("strided" if tag & 2 else "contiguous")
        passes[nm] = {"ms": sum(v) / len(v), "launches": len(v) // reps,
                      "kernel": f"fft_tma_kernel<{tag & 1}, {(tag >> 1) & 1}, 1, {(tag >> 2) & 1}>", "tensor_core_top_stage": bool(tag & 4)}
    leaf = layers.data_ptr() + 32 * ((1 << ext) - 1)
    t_leaf = timed(lambda: chk(L.tstwo_merkle_commit_layer(ext, 0, ep, n_cols, leaf)))

    def nodes():
        prev = leaf
        for k in range(ext - 1, -1, -1):
            cur = layers.data_ptr() + 32 * ((1 << k) - 1)
            chk(L.tstwo_merkle_commit_layer(k, prev, None, 0, cur))
            prev = cur
    t_nodes = timed(nodes)
    b_int, b_ev = 8 * n_cols * n, (4 + 4 * (1 << blow)) * n_cols * n
    b_leaf, b_node = (4 * n_cols + 32) * ne, 96 * (ne - 1)
    mk = lambda ms, by, launches: {"ms": ms, "bytes": by, "gbs": by / ms / 1e6, "launches": launches}
    return {"interpolate (2 fft launches)": mk(t_int, b_int, 2), "evaluate (2 fft launches)": mk(t_ev, b_ev, 2),
            "merkle_leaf_layer": mk(t_leaf, b_leaf, 1), "merkle_node_layers": mk(t_nodes, b_node, ext), "passes": passes}


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    elif a.single_process:
        run_single_process(a)
    else:
        run_cuda(a)

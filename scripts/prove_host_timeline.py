"""Where prove()'s wall clock goes on the host: time inside C-ABI calls (enqueue + waits for the device) against time in the Python
mirror between them (during which the device idles once the previous call has synchronised).  usage: ... [log=23]"""
import os, sys, time, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import tstwo_b200 as tb
from tstwo_b200 import _lib, air, prover
from tstwo_b200.channel import Blake2sChannel
P = 2147483647
log = int(sys.argv[1]) if len(sys.argv) > 1 else 23
n_cols, blow = 100, 1
L = _lib.init(0)
n = 1 << log
trace = torch.empty((n_cols, n), dtype=torch.int32, device="cuda")
a = torch.ones(n, dtype=torch.int64, device="cuda"); b = torch.arange(n, dtype=torch.int64, device="cuda")
trace[0], trace[1] = a.to(torch.int32), b.to(torch.int32)
for k in range(2, n_cols):
    a, b = b, (a * a % P + b * b % P) % P
    trace[k] = b.to(torch.int32)
del a, b
tw = tb.CudaBackend.precomputeTwiddles(tb.CanonicCoset.new(log + 1 + blow).half_coset())
dom = tb.CanonicCoset.new(log).circleDomain()
cols = [tb.CudaCircleEvaluation(dom, tb.CudaColumn.view(trace.data_ptr() + 4 * (c << log), n, trace)) for c in range(n_cols)]

calls = []
class Timed:
    def __init__(self, lib): object.__setattr__(self, "_lib", lib)
    def __getattr__(self, name):
        fn = getattr(self._lib, name)
        if not name.startswith("tstwo_"): return fn
        def wrapped(*a):
            t0 = time.perf_counter(); r = fn(*a); calls.append((name, t0, time.perf_counter())); return r
        return wrapped
real = _lib._lib
for rep in range(3):
    channel = Blake2sChannel.create()
    scheme = tb.CommitmentSchemeProver(tb.PcsConfig(10, tb.FriConfig(0, blow, 70)), tw)
    bld = scheme.tree_builder(); bld.extend_evals([]); bld.commit(channel)
    bld = scheme.tree_builder(); bld.extend_evals(cols); bld.commit(channel)
    real.tstwo_sync()
    if rep == 2: _lib._lib = Timed(real); calls.clear()
    t0 = time.perf_counter()
    proof = prover.prove([air.WideFibonacciComponent(air.WideFibonacciEval.create(log, n_cols))], channel, scheme)
    real.tstwo_sync(); t1 = time.perf_counter()
    _lib._lib = real
    print(f"pass {rep}: prove() {1e3 * (t1 - t0):.2f} ms")
    del proof, scheme
inside = sum(e - s for _, s, e in calls)
print(f"{len(calls)} C-ABI calls, {1e3 * inside:.2f} ms inside them, {1e3 * (t1 - t0 - inside):.2f} ms in Python between them")
agg = collections.OrderedDict()
for nm, s, e in calls:
    v = agg.setdefault(nm, [0, 0.0]); v[0] += 1; v[1] += e - s
for nm, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:14]:
    print(f"  {nm:40s} {c:4d} calls {1e3 * t:8.3f} ms")
gaps = sorted(((calls[i + 1][1] - calls[i][2], calls[i][0], calls[i + 1][0]) for i in range(len(calls) - 1)), reverse=True)[:12]
print("largest gaps between calls (ms, after, before):")
for g, a_, b_ in gaps: print(f"  {1e3 * g:7.3f}  {a_} -> {b_}")

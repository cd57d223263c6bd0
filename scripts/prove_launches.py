"""One prove() of BASELINE config 5 (wide-Fibonacci, 100 columns x 2^log, blow-up 2) after a warm-up pass, for an ncu launch list:
ncu --metrics gpu__time_duration.sum --csv --log-file gpurun_out/prove_launches.csv python scripts/prove_launches.py [log=23]
(the script prints the number of kernels launched before the measured pass so that the list can be cut there)."""
import os, sys, time
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
for rep in range(2):
    channel = Blake2sChannel.create()
    scheme = tb.CommitmentSchemeProver(tb.PcsConfig(10, tb.FriConfig(0, blow, 70)), tw)
    bld = scheme.tree_builder(); bld.extend_evals([]); bld.commit(channel)
    bld = scheme.tree_builder(); bld.extend_evals(cols); bld.commit(channel)
    L.tstwo_sync(); l0 = L.tstwo_launch_count(); t0 = time.perf_counter()
    proof = prover.prove([air.WideFibonacciComponent(air.WideFibonacciEval.create(log, n_cols))], channel, scheme)
    L.tstwo_sync()
    print(f"pass {rep}: prove() {1e3 * (time.perf_counter() - t0):.2f} ms, {L.tstwo_launch_count() - l0} library launches, {l0} before it", flush=True)
    del proof, scheme

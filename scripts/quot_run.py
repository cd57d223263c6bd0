"""Runs accumulate_quotients on 256 LDE columns of 2^22 rows a few times (for ncu / timing)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import tstwo_b200 as tb
from tstwo_b200 import _lib
P = 2147483647
L = _lib.init(0)
ncols, qlog = 256, int(sys.argv[1]) if len(sys.argv) > 1 else 22
rng = np.random.default_rng(1)
block = torch.randint(0, P, (ncols, 1 << qlog), dtype=torch.int32, device="cuda")
qdom = tb.CanonicCoset.new(qlog).circleDomain()
evs = [tb.CudaCircleEvaluation(qdom, tb.CudaColumn.view(block.data_ptr() + 4 * (c << qlog), 1 << qlog, block)) for c in range(ncols)]
pt = ((1, 0, 478637715, 513582971), (992285211, 649143431, 740191619, 1186584352))
batches = [(pt, [(c, [int(x) for x in rng.integers(0, P, 4)]) for c in range(ncols)])]
alpha = [1, 2, 3, 4]
mode = int(sys.argv[2]) if len(sys.argv) > 2 else 0
L.tstwo_debug_set_quot_mode(mode)
for _ in range(4):
    t0 = time.perf_counter(); tb.accumulate_quotients(qdom, evs, alpha, batches, 1, ts_denominators=False); L.tstwo_sync()
    print(f"{(time.perf_counter() - t0) * 1e3:.3f} ms (API level)")
if mode == 0:   # the two kernel forms must agree bit for bit
    outs = []
    for m in (1, 2):
        L.tstwo_debug_set_quot_mode(m)
        o = tb.accumulate_quotients(qdom, evs, alpha, batches, 1, ts_denominators=False)
        outs.append([c.copy() for c in o.values.to_cpu()])
    L.tstwo_debug_set_quot_mode(0)
    print("integer-pipe and tensor-core kernels agree:", all((a == b).all() for a, b in zip(*outs)))

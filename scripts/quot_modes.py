"""Times the two forms of the DEEP-quotient kernel (debug mode 1: integer pipes, 2: INT8 tensor cores) with CUDA events."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import tstwo_b200 as tb
from tstwo_b200 import _lib
P = 2147483647
L = _lib.init(0)
ncols, qlog = 256, 22
rng = np.random.default_rng(1)
block = torch.randint(0, P, (ncols, 1 << qlog), dtype=torch.int32, device="cuda")
qdom = tb.CanonicCoset.new(qlog).circleDomain()
evs = [tb.CudaCircleEvaluation(qdom, tb.CudaColumn.view(block.data_ptr() + 4 * (c << qlog), 1 << qlog, block)) for c in range(ncols)]
pt = ((1, 0, 478637715, 513582971), (992285211, 649143431, 740191619, 1186584352))
batches = [(pt, [(c, [int(x) for x in rng.integers(0, P, 4)]) for c in range(ncols)])]
for mode in (1, 2):
    L.tstwo_debug_set_quot_mode(mode)
    ts = []
    for _ in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        tb.accumulate_quotients(qdom, evs, [1, 2, 3, 4], batches, 1, ts_denominators=False)
        e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    print(f"mode {mode} ({'integer pipes' if mode == 1 else 'tensor cores'}): API call {min(ts):.3f} ms")

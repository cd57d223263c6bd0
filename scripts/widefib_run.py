"""Times tstwo_widefib_constraints on BASELINE config 2's shape (256 columns, trace log 20 -> evaluation domain 2^21) in both
kernel forms (debug mode 1: integer pipes, 2: INT8 tensor cores), CUDA events around the API call."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import json
import numpy as np, torch
import tstwo_b200 as tb
from tstwo_b200 import _lib, air
P = 2147483647
L = _lib.init(0)
n_cols, tlog = 256, int(sys.argv[1]) if len(sys.argv) > 1 else 20
elog = tlog + 1
peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else {}
hbm = float(peaks.get("hbm_gbs", 6455.0))
rng = np.random.default_rng(2)
block = torch.randint(0, P, (n_cols, 1 << elog), dtype=torch.int32, device="cuda")
cols = [tb.CudaColumn.view(block.data_ptr() + 4 * (c << elog), 1 << elog, block) for c in range(n_cols)]
coeffs = [[int(x) for x in rng.integers(0, P, 4)] for _ in range(n_cols - 2)]
den = [int(x) for x in rng.integers(1, P, 2)]
outs = {}
for mode in (1, 2):
    L.tstwo_debug_set_quot_mode(mode)
    out = tb.SecureColumnByCoords.zeros(1 << elog)
    ts = []
    for _ in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        air.widefib_constraints(tlog, elog, cols, coeffs, den, out)
        e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    byt = (n_cols + 8) * 4 * (1 << elog)
    print(f"mode {mode} ({'integer pipes' if mode == 1 else 'tensor cores'}): API call {min(ts):.3f} ms, {byt / min(ts) / 1e6:.0f} GB/s, "
          f"{100 * byt / min(ts) / 1e6 / hbm:.1f} % of HBM peak ({hbm:.0f} GB/s)")
    outs[mode] = [c.copy() for c in out.to_cpu()]
L.tstwo_debug_set_quot_mode(0)
print("both forms agree:", all((a == b).all() for a, b in zip(outs[1], outs[2])))

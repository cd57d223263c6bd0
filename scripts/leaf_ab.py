"""A/B timing of the leaf-layer kernels at the headline size (256 columns x 2^23 rows): bulk-copy-fed vs per-thread loads.
Run on the GPU box: python scripts/leaf_ab.py [n_cols] [log_rows]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tstwo_b200 import _lib

n_cols = int(sys.argv[1]) if len(sys.argv) > 1 else 256
log = int(sys.argv[2]) if len(sys.argv) > 2 else 23
L = _lib.load(); _lib.init(0)
ne = 1 << log
g = torch.Generator(device="cuda").manual_seed(1)
evals = torch.randint(0, 2**31 - 1, (n_cols, ne), dtype=torch.int32, device="cuda", generator=g)
if os.environ.get("LEAF_CONST"):
    evals.fill_(0x05050505)
out = [torch.empty((ne, 8), dtype=torch.int32, device="cuda") for _ in range(2)]
ep = _lib.dptr_array([evals.data_ptr() + 4 * ne * c for c in range(n_cols)])
import ctypes as C
stream_ptr = C.c_void_p()
s = torch.cuda.current_stream()
def run(mode, dst):
    L.tstwo_debug_set_leaf_mode(mode)
    _lib.check(L.tstwo_merkle_commit_layer(log, 0, ep, n_cols, dst.data_ptr()))
    assert L.tstwo_debug_last_leaf_mode() == mode, (mode, L.tstwo_debug_last_leaf_mode())
outs = {}
for mode in (1, 0):
    outs[mode] = out[1 if mode else 0]
    run(mode, outs[mode]); L.tstwo_sync(); torch.cuda.synchronize()
    import time
    ts = []
    for _ in range(5):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        run(mode, outs[mode]); L.tstwo_sync()
        ts.append((time.perf_counter() - t0) * 1e3)
    n_cmp = ne * ((n_cols + 15) // 16)
    best = min(ts)
    print(f"mode {mode}: {['per-thread loads', 'bulk-copy-fed'][mode]:18s} best {best:.3f} ms  median {sorted(ts)[2]:.3f} ms  "
          f"{n_cmp / best / 1e6:.2f} G compress/s  {best * 1e-3 * 1.965e9 * 148 * 4 / (n_cmp / 32):.0f} clk per warp-compress per SMSP")
print("equal:", bool((out[0] == out[1]).all()))

"""Per-pass CUDA-event times of interpolate + evaluate (tstwo_lde_commit without the tree) for a given shape.
usage: python scripts/pass_times.py n_cols log_size [log_blowup=1]"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tstwo_b200 import _lib
P = 2147483647
n_cols, log = int(sys.argv[1]), int(sys.argv[2])
blow = int(sys.argv[3]) if len(sys.argv) > 3 else 1
ext = log + blow
L = _lib.init(0); chk = _lib.check
trace = torch.randint(0, P, (n_cols, 1 << log), dtype=torch.int32, device="cuda")
coeffs = torch.empty((n_cols, 1 << log), dtype=torch.int32, device="cuda")
evals = torch.empty((n_cols, 1 << ext), dtype=torch.int32, device="cuda")
tw = C.c_uint64(); chk(L.tstwo_twiddles_precompute(1 << (31 - (ext - 1) - 2), ext - 1, C.byref(tw)))
root = (C.c_uint8 * 32)()
run = lambda: chk(L.tstwo_lde_commit(trace.data_ptr(), n_cols, log, blow, tw.value, coeffs.data_ptr(), evals.data_ptr(), 0, root, 0))
for _ in range(3): run()
L.tstwo_sync()
L.tstwo_debug_pass_timing(1)
for _ in range(5): run()
L.tstwo_sync()
tags, ms = (C.c_int * 256)(), (C.c_float * 256)()
k = L.tstwo_debug_pass_times(256, tags, ms)
per = k // 5
print(f"{n_cols} x 2^{log}, blow-up 2^{blow}: {per} passes per LDE")
tot = 0
for i in range(per):
    t = sum(ms[i + per * r] for r in range(5)) / 5
    tot += t
    tag = tags[i]
    print(f"  pass {i}: {'inverse' if tag & 1 else 'forward'} {'strided' if tag & 2 else 'contiguous'}{' (tensor-core top stage)' if tag & 4 else ''}  {t:7.3f} ms")
print(f"  total {tot:.3f} ms = {n_cols * (1 << log) / tot / 1e6:.1f} G elements/s")

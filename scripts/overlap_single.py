"""Does the two-stream overlap of tstwo_lde_columns_to_peers_overlapped (ALU-bound contiguous pass of column group k beside the
DRAM-heavier passes of group k + 1, SMs partitioned) pay on ONE GPU?  n_shards = 1: the "peer" buffer is local.
usage: python scripts/overlap_single.py [g,k ...]"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tstwo_b200 import _lib
P = 2147483647
L = _lib.init(0)
chk = _lib.check
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
chk(L.tstwo_set_stream(C.c_void_p(stream.cuda_stream)))
n_cols, log, blow = 256, 22, 1
ext = log + blow
g = torch.Generator(device="cuda"); g.manual_seed(7)
trace = torch.randint(0, P, (n_cols, 1 << log), dtype=torch.int32, device="cuda", generator=g)
coeffs = torch.empty((n_cols, 1 << log), dtype=torch.int32, device="cuda")
work = torch.empty((n_cols, 1 << ext), dtype=torch.int32, device="cuda")
evals = torch.empty((n_cols, 1 << ext), dtype=torch.int32, device="cuda")
ref = torch.empty((n_cols, 1 << ext), dtype=torch.int32, device="cuda")
tw = C.c_uint64()
chk(L.tstwo_twiddles_precompute(1 << (31 - (ext - 1) - 2), ext - 1, C.byref(tw)))
bases = _lib.dptr_array([evals.data_ptr()])
cfgs = [tuple(int(x) for x in a.split(",")) for a in sys.argv[1:]] or [(1, 0), (2, 62), (4, 62), (4, 50), (4, 74), (8, 62), (1, 0)]
first = None
for cfg in cfgs:
    def run():
        chk(L.tstwo_lde_columns_to_peers_overlapped(trace.data_ptr(), n_cols, log, blow, tw.value, coeffs.data_ptr(), work.data_ptr(), bases, 1, 0, 0,
                                                    cfg[0], cfg[1]))
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(10):
        run()
    b.record(stream); torch.cuda.synchronize()
    if first is None:
        ref.copy_(evals); first = True
    same = bool((ref == evals).all())
    print(f"groups={cfg[0]:2d} side_sms={cfg[1]:3d}  LDE {a.elapsed_time(b) / 10:7.3f} ms  same evaluations: {same}", flush=True)

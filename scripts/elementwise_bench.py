"""Times the remaining elementwise entry points at large sizes against the measured HBM peak (CUDA events around the C-ABI call,
best of 10): bit_reverse, accumulate, fri_decompose, eval_at_point."""
import ctypes as C, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import tstwo_b200 as tb
from tstwo_b200 import _lib
from tstwo_b200._lib import check, dptr_array, u32_array
P = 2147483647
L = _lib.init(0)
pk = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "MEASURED_PEAKS.json")
hbm = float(json.load(open(pk)).get("hbm_gbs", 6455.0)) if os.path.exists(pk) else 6455.0
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
check(L.tstwo_set_stream(C.c_void_p(stream.cuda_stream)))

def timed(fn, reps=10):
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record(stream); fn(); e1.record(stream); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best

def report(name, ms, nbytes):
    print(f"{name:58s} {ms * 1e3:9.1f} us  {nbytes / ms / 1e6:7.0f} GB/s  {100 * nbytes / ms / 1e6 / hbm:5.1f} % of HBM peak")

print(f"# HBM peak {hbm:.0f} GB/s")
for log in (20, 23, 26):
    col = torch.randint(0, P, (1 << log,), dtype=torch.int32, device="cuda")
    report(f"bit_reverse 2^{log}", timed(lambda: check(L.tstwo_bit_reverse(col.data_ptr(), 1 << log))), 8 << log)
for log in (20, 24):
    a = torch.randint(0, P, (4, 1 << log), dtype=torch.int32, device="cuda")
    b = torch.randint(0, P, (4, 1 << log), dtype=torch.int32, device="cuda")
    pa, pb = dptr_array([a[k].data_ptr() for k in range(4)]), dptr_array([b[k].data_ptr() for k in range(4)])
    report(f"accumulate 4 x 2^{log}", timed(lambda: check(L.tstwo_accumulate(pa, pb, 1 << log))), 3 * 16 << log)
    g = torch.empty((4, 1 << log), dtype=torch.int32, device="cuda")
    pg = dptr_array([g[k].data_ptr() for k in range(4)])
    lam = (C.c_uint32 * 4)()
    report(f"fri_decompose 4 x 2^{log} (sum pass + apply pass)", timed(lambda: check(L.tstwo_fri_decompose(pa, log, pg, lam))), 3 * 16 << log)
for log in (16, 22, 24):
    co = torch.randint(0, P, (1 << log,), dtype=torch.int32, device="cuda")
    out = (C.c_uint32 * 4)()
    px, py = u32_array([1, 0, 478637715, 513582971]), u32_array([992285211, 649143431, 740191619, 1186584352])
    report(f"eval_at_point 2^{log} coefficients", timed(lambda: check(L.tstwo_eval_at_point(co.data_ptr(), log, px, py, out))), 4 << log)
for log, cnt in ((20, 100), (23, 100)):
    block = torch.randint(0, P, (cnt, 1 << log), dtype=torch.int32, device="cuda")
    ptrs = dptr_array([block[k].data_ptr() for k in range(cnt)])
    res = (C.c_uint32 * (4 * cnt))()
    px, py = u32_array([1, 0, 478637715, 513582971]), u32_array([992285211, 649143431, 740191619, 1186584352])
    report(f"eval_at_point_many {cnt} x 2^{log} coefficients", timed(lambda: check(L.tstwo_eval_at_point_many(ptrs, cnt, log, px, py, res)), 5),
           cnt * (4 << log))          # every coefficient read once; the partial sums are n / 2048 QM31 values

"""Determinism soak of the shipped path: the headline job (256 x 2^22, blow-up 2) REPS times, the root checked every time; then
round trips at log 23 / 24 / 25 (three-pass transforms, partial windows) and the FRI commit of config 4, each repeated.
usage: python scripts/soak.py [reps=40]"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import tstwo_b200 as tb
from tstwo_b200 import _lib
from tstwo_b200.channel import Blake2sChannel
from tstwo_b200.fri import FriProver
P = 2147483647
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 40
L = _lib.init(0); chk = _lib.check
n_cols, log, blow = 256, 22, 1
ext = log + blow
g = torch.Generator(device="cuda"); g.manual_seed(99)
trace = torch.randint(0, P, (n_cols, 1 << log), dtype=torch.int32, device="cuda", generator=g)
coeffs = torch.empty((n_cols, 1 << log), dtype=torch.int32, device="cuda")
evals = torch.empty((n_cols, 1 << ext), dtype=torch.int32, device="cuda")
layers = torch.empty(((2 << ext) - 1) * 8, dtype=torch.int32, device="cuda")
tw = C.c_uint64(); chk(L.tstwo_twiddles_precompute(1 << (31 - (ext - 1) - 2), ext - 1, C.byref(tw)))
roots = set()
for _ in range(reps):
    root = (C.c_uint8 * 32)()
    chk(L.tstwo_lde_commit(trace.data_ptr(), n_cols, log, blow, tw.value, coeffs.data_ptr(), evals.data_ptr(), layers.data_ptr(), root, 0))
    roots.add(bytes(root))
print(f"headline job x {reps}: {len(roots)} distinct root(s) {next(iter(roots)).hex()[:16]}", flush=True)
assert len(roots) == 1
del trace, coeffs, evals, layers
torch.cuda.empty_cache(); L.tstwo_mem_trim()
for lg in (23, 24, 25):
    twd = tb.CudaBackend.precomputeTwiddles(tb.Coset.half_odds(lg - 1))
    vals = np.random.default_rng(lg).integers(0, P, 1 << lg, dtype=np.uint32)
    dom = tb.CanonicCoset.new(lg).circleDomain()
    ref = None
    for it in range(max(4, reps // 4)):
        poly = tb.CudaCircleEvaluation.new(dom, vals).interpolateWithTwiddles(twd)
        c = poly.coeffs.toCpu()
        ref = c if ref is None else ref
        back = poly.evaluateWithTwiddles(dom, twd).values.toCpu()
        assert (c == ref).all() and (back == vals).all(), f"log {lg} iteration {it}"
    print(f"round trip log {lg} x {max(4, reps // 4)}: identical", flush=True)
    del twd
lg = 22
dom = tb.CanonicCoset.new(lg).circleDomain()
twd = tb.CudaBackend.precomputeTwiddles(dom.halfCoset)
rng = np.random.default_rng(4)
coords = [tb.CudaCirclePoly.new(rng.integers(0, P, 1 << (lg - 1), dtype=np.uint32)).evaluateWithTwiddles(dom, twd).values.toCpu() for _ in range(4)]
col = tb.SecureEvaluation(dom, tb.SecureColumnByCoords.from_cpu(coords))
digests = set()
for _ in range(max(4, reps // 4)):
    ch = Blake2sChannel.create()
    pr = FriProver.commit(ch, tb.FriConfig(0, 1, 20), [col], twd)
    pr.decommit(ch)
    digests.add(ch.digest())
print(f"FRI commit + decommit log {lg} x {max(4, reps // 4)}: {len(digests)} distinct transcript(s)")
assert len(digests) == 1
print("soak OK")

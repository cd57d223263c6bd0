"""Repeat evaluate(interpolate(v)) at a large size in both pass families and report where a mismatch lands."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import tstwo_b200 as tb
from tstwo_b200 import _lib
P = 2147483647
log = int(sys.argv[1]) if len(sys.argv) > 1 else 24
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
L = _lib.init(0)
tw = tb.CudaBackend.precomputeTwiddles(tb.Coset.half_odds(log - 1))
vals = np.random.default_rng(log).integers(0, P, 1 << log, dtype=np.uint32)
dom = tb.CanonicCoset.new(log).circleDomain()
ref_c = None
for it in range(reps):
    for mode in (2, 1):
        L.tstwo_debug_set_fft_stream(mode)
        poly = tb.CudaCircleEvaluation.new(dom, vals).interpolateWithTwiddles(tw)
        c = poly.coeffs.toCpu()
        if ref_c is None: ref_c = c
        badc = np.nonzero(c != ref_c)[0]
        back = poly.evaluateWithTwiddles(dom, tw).values.toCpu()
        bad = np.nonzero(back != vals)[0]
        if len(bad) or len(badc):
            def runs(ix):
                if not len(ix): return "-"
                return f"{len(ix)} in [{ix[0]:#x}, {ix[-1]:#x}] first run ends {ix[np.nonzero(np.diff(ix) > 1)[0][0]] if (np.diff(ix) > 1).any() else ix[-1]:#x}"
            print(f"iter {it} mode {mode}: coeff mismatches {runs(badc)}; eval mismatches {runs(bad)}", flush=True)
L.tstwo_debug_set_fft_stream(2)
print("done", log, reps)

"""One FriProver.commit of BASELINE config 4 (one QM31 column on CanonicCoset(24)) for an ncu launch list:
ncu --metrics gpu__time_duration.sum --csv --log-file gpurun_out/fri_launches.csv python scripts/fri_commit_profile.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import tstwo_b200 as tb
from tstwo_b200 import _lib
from tstwo_b200.channel import Blake2sChannel
from tstwo_b200.fri import FriProver
P = 2147483647
log = int(sys.argv[1]) if len(sys.argv) > 1 else 24
L = _lib.init(0)
rng = np.random.default_rng(0x5EED0004)
dom = tb.CanonicCoset.new(log).circleDomain()
tw = tb.CudaBackend.precomputeTwiddles(dom.halfCoset)
coords = [tb.CudaCirclePoly.new(rng.integers(0, P, 1 << (log - 1), dtype=np.uint32)).evaluateWithTwiddles(dom, tw).values for _ in range(4)]
sc = tb.SecureColumnByCoords.uninitialized(1 << log)      # one block, as the prover's own steps produce a secure column
for k in range(4):
    L.tstwo_copy(sc.columns[k].ptr, coords[k].ptr, 4 << log)
col = tb.SecureEvaluation(dom, sc)
cfg = tb.FriConfig(0, 1, 30)
for rep in range(3):
    L.tstwo_sync(); t0 = time.perf_counter()
    pr = FriProver.commit(Blake2sChannel.create(), cfg, [col], tw)
    L.tstwo_sync(); print("FriProver.commit %.3f ms" % ((time.perf_counter() - t0) * 1e3))
    del pr

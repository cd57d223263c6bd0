"""BASELINE config 4 on the device: FRI commit of one QM31 column on CanonicCoset(24) (fold_circle_into_line, then fold_line down
to the last layer), kernel by kernel against the HBM roofline, then the whole FriProver.commit (folds + one Merkle tree per
layer + transcript).  Also the DEEP-quotient kernel on 256 LDE columns.  usage: python scripts/fri_bench.py [log_size]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import tstwo_b200 as tb
from tstwo_b200 import _lib
from tstwo_b200.channel import Blake2sChannel
from tstwo_b200.fri import FriProver

P = 2147483647
log = int(sys.argv[1]) if len(sys.argv) > 1 else 24
L = _lib.init(0)
try:
    peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json"))).get("hbm_gbs", 6455.3)
except Exception:
    peak = 6455.3


def timed(fn, reps=20):
    fn(); L.tstwo_sync()
    best = 1e9
    for _ in range(3):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        L.tstwo_sync(); best = min(best, (time.perf_counter() - t0) / reps)
    return best * 1e3


rng = np.random.default_rng(0x5eed0004)
dom = tb.CanonicCoset.new(log).circleDomain()
src = tb.SecureEvaluation(dom, tb.SecureColumnByCoords.from_cpu([rng.integers(0, P, 1 << log, dtype=np.uint32) for _ in range(4)]))
tw = tb.CudaBackend.precomputeTwiddles(tb.CanonicCoset.new(log).half_coset())
alpha = [int(x) for x in rng.integers(0, P, 4)]
print(f"# FRI folds of one QM31 column on CanonicCoset({log}) (BASELINE config 4), HBM peak {peak:.0f} GB/s")
line = tb.LineEvaluation.newZero(tb.LineDomain.new(tb.Coset.half_odds(log - 1)))
n = 1 << log
t = timed(lambda: tb.fold_circle_into_line(line, src, alpha, tw))
by = 16 * n + 2 * 16 * (n // 2)
print(f"fold_circle_into_line 2^{log} -> 2^{log-1}: {t*1e3:8.1f} us  {by/t/1e6:7.0f} GB/s  {by/t/1e6/peak*100:5.1f} % of HBM peak")
total_t, total_b = t, by
cur = line
while cur.len() > 2:
    m = cur.len()
    t = timed(lambda: tb.fold_line(cur, alpha, tw), reps=20 if m > 1 << 16 else 50)
    by = 16 * m + 16 * (m // 2) + 4 * (m // 2)
    if m >= 1 << 12:
        print(f"fold_line 2^{m.bit_length()-1} -> 2^{m.bit_length()-2}: {t*1e3:8.1f} us  {by/t/1e6:7.0f} GB/s  {by/t/1e6/peak*100:5.1f} % of HBM peak")
    total_t += t; total_b += by
    cur = tb.fold_line(cur, alpha, tw)
print(f"all folds: {total_t*1e3:.1f} us for {total_b/2**30:.2f} GiB algorithmic => {total_b/total_t/1e6:.0f} GB/s ({total_b/total_t/1e6/peak*100:.1f} % of HBM peak; "
      f"{(1<<log)/total_t/1e6:.2f} G QM31 elements/s)")

# the whole commit phase: folds + Merkle tree per layer + transcript (low-degree input so that the last layer accepts)
deg = log - 1
orc_free_cols = []
root = tb.CanonicCoset.new(log).half_coset()
coords = []
for k in range(4):
    c = np.zeros(1 << deg, dtype=np.uint32); c[: 1 << deg] = rng.integers(0, P, 1 << deg, dtype=np.uint32)
    coords.append(tb.CudaCirclePoly.new(c).evaluateWithTwiddles(dom, tw).values)
sc = tb.SecureColumnByCoords.uninitialized(1 << log)      # one block, as the prover's own steps produce a secure column
for k in range(4):
    L.tstwo_copy(sc.columns[k].ptr, coords[k].ptr, 4 << log)
col = tb.SecureEvaluation(dom, sc)
cfg = tb.FriConfig(0, 1, 30)
def commit():
    return FriProver.commit(Blake2sChannel.create(), cfg, [col], tw)
commit(); L.tstwo_sync()
t0 = time.perf_counter(); pr = commit(); L.tstwo_sync(); t_commit = time.perf_counter() - t0
t0 = time.perf_counter(); proof, _ = pr.decommit(Blake2sChannel.create()); t_dec = time.perf_counter() - t0
print(f"FriProver.commit (first-layer tree, {len(pr.inner_layers)} inner layers each folded + committed, last layer poly): {t_commit*1e3:.2f} ms; "
      f"decommit of {cfg.n_queries} queries: {t_dec*1e3:.2f} ms")

# DEEP quotients: 256 LDE columns of 2^22 rows sampled at one point
ncols, qlog = 256, 22
block = torch.randint(0, P, (ncols, 1 << qlog), dtype=torch.int32, device="cuda")
qdom = tb.CanonicCoset.new(qlog).circleDomain()
evs = [tb.CudaCircleEvaluation(qdom, tb.CudaColumn.view(block.data_ptr() + 4 * (c << qlog), 1 << qlog, block)) for c in range(ncols)]
pt = ((1, 0, 478637715, 513582971), (992285211, 649143431, 740191619, 1186584352))
batches = [(pt, [(c, [int(x) for x in rng.integers(0, P, 4)]) for c in range(ncols)])]
t = timed(lambda: tb.accumulate_quotients(qdom, evs, alpha, batches, 1, ts_denominators=False), reps=5)
by = 4 * ncols * (1 << qlog) + 16 * (1 << qlog)
print(f"accumulate_quotients {ncols} cols x 2^{qlog}, 1 sample point: {t:.3f} ms  {by/t/1e6:.0f} GB/s  {by/t/1e6/peak*100:.1f} % of HBM peak")

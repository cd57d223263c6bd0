"""cProfile of prove() on the device backend (host time: which Python frames / C-ABI calls the wall clock goes to).
usage: python scripts/prove_profile.py [log_n_rows=20] [n_cols=100]"""
import cProfile, os, pstats, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tstwo_b200 as tb
from tstwo_b200 import _lib, air, prover
from tstwo_b200.channel import Blake2sChannel

log = int(sys.argv[1]) if len(sys.argv) > 1 else 20
n_cols = int(sys.argv[2]) if len(sys.argv) > 2 else 100
L = _lib.init(0)
trace = air.generate_wide_fibonacci_trace(log, n_cols)
tw = tb.CudaBackend.precomputeTwiddles(tb.CanonicCoset.new(log + 2).half_coset())
dom = tb.CanonicCoset.new(log).circleDomain()
cols = [tb.CudaCircleEvaluation.new(dom, c) for c in trace]

def once(profile=None):
    channel = Blake2sChannel.create()
    scheme = tb.CommitmentSchemeProver(tb.PcsConfig(10, tb.FriConfig(0, 1, 70)), tw)
    b = scheme.tree_builder(); b.extend_evals([]); b.commit(channel)
    b = scheme.tree_builder(); b.extend_evals(cols); b.commit(channel)
    comp = air.WideFibonacciComponent(air.WideFibonacciEval.create(log, n_cols))
    L.tstwo_sync(); t = time.perf_counter()
    if profile: profile.enable()
    proof = prover.prove([comp], channel, scheme)
    if profile: profile.disable()
    L.tstwo_sync()
    return time.perf_counter() - t

once(); once()
print("prove() %.1f ms (unprofiled)" % (1e3 * once()))
pr = cProfile.Profile()
print("prove() %.1f ms (profiled)" % (1e3 * once(pr)))
st = pstats.Stats(pr); st.sort_stats("cumulative").print_stats(45)
st.sort_stats("tottime").print_stats(25)

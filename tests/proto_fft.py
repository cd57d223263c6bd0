"""numpy prototype of the tiled circle-FFT pass decomposition used by tstwo_b200/csrc/fft_fast.cu.
Validates the index math (window/tile/twiddle indices, coset handling, active-layer masks) against the oracle
before it is transcribed to CUDA.  Not part of the product or the tests."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle as orc

P = np.uint64(2**31 - 1)
def madd(a, b): return ((a.astype(np.uint64) + b) % P).astype(np.uint32)
def msub(a, b): return ((a.astype(np.uint64) + P - b) % P).astype(np.uint32)
def mmul(a, b): return ((a.astype(np.uint64) * b.astype(np.uint64)) % P).astype(np.uint32)

W0, W1 = 12, 10  # window bits: contiguous pass, strided passes

def layer_table_offset(tree_log, n, i):
    """FFT layer i>=1 of a domain of log n: T_i has 2^(n-1-i) entries at offset 2^tree_log - 2^(n-i)."""
    return (1 << tree_log) - (1 << (n - i))

def twiddle(tree, tree_log, n, i, h):
    """twiddle for layer i (0 = circle), block index h = idx >> (i+1), domain log n."""
    if i >= 1:
        return tree[layer_table_offset(tree_log, n, i) + h]
    t1 = tree[layer_table_offset(tree_log, n, 1) + ((h >> 1) ^ 1)]
    sign = ((h ^ (h >> 1)) & 1).astype(bool)
    return np.where(sign, (P - t1.astype(np.uint64)) % P, t1).astype(np.uint32)

def run_pass(src, dst, n, src_mask, lo, j_min, j_max, tree, tree_log, inverse, W):
    """One pass over the window of row bits [lo, lo+W).  src index = idx & src_mask (coset/zero-extension trick).
    Processes layers lo+j for j in [j_min, j_max)."""
    N = 1 << n
    idx = np.arange(N, dtype=np.uint64)
    row = (idx >> lo) & ((1 << W) - 1)
    tile_hi = idx >> (lo + W)
    v = src[idx & src_mask].copy()
    js = range(j_min, j_max) if inverse else range(j_max - 1, j_min - 1, -1)
    for j in js:
        i = lo + j
        h = (tile_hi << (W - 1 - j)) | (row >> (j + 1))      # == idx >> (i+1)
        assert (h == (idx >> (i + 1))).all()
        t = twiddle(tree, tree_log, n, i, h)
        is_lo = ((row >> j) & 1) == 0
        partner = idx ^ (1 << i)
        a = np.where(is_lo, v, v[partner]); b = np.where(is_lo, v[partner], v)
        if inverse:
            out = np.where(is_lo, madd(a, b), mmul(msub(a, b), t))
        else:
            m = mmul(b, t)
            out = np.where(is_lo, madd(a, m), msub(a, m))
        v = out
    dst[:] = v

def plan(L):
    """passes as (lo, j_min, j_max, W) in ascending layer order"""
    if L < 14: return None
    out = [(0, 0, W0, W0)]
    done = W0
    while L - done > W1:
        out.append((done, 0, W1, W1)); done += W1
    out.append((L - W1, done - (L - W1), W1, W1))
    return out

def interpolate(vals, n, tree_itw, tree_log):
    buf = vals.copy()
    for (lo, a, b, W) in plan(n):
        out = np.empty_like(buf)
        run_pass(buf, out, n, (1 << n) - 1, lo, a, b, tree_itw, tree_log, True, W)
        buf = out
    inv_n = pow(1 << n, 2**31 - 3, 2**31 - 1)
    return mmul(buf, np.full(1, inv_n, dtype=np.uint32))

def evaluate(coeffs, L, n, tree_tw, tree_log):
    src = coeffs
    mask = (1 << L) - 1
    for (lo, a, b, W) in reversed(plan(L)):
        out = np.empty(1 << n, dtype=np.uint32)
        run_pass(src, out, n, mask, lo, a, b, tree_tw, tree_log, False, W)
        src, mask = out, (1 << n) - 1
    return src

if __name__ == "__main__":
    rng = np.random.default_rng(0)
    for (L, blow, extra) in [(14, 0, 0), (14, 1, 0), (15, 2, 1), (16, 1, 0), (18, 1, 2), (22, 0, 0), (23, 0, 0)]:
        n = L + blow
        tree_log = n - 1 + extra
        init = orc.half_odds_index(tree_log)
        tw, itw = orc.precompute_twiddles(init, tree_log)
        vals = rng.integers(0, 2**31 - 1, 1 << L, dtype=np.uint32)
        h_small = orc.canonic_half_coset(L)
        exp_c = orc.interpolate(vals, h_small[0], (init, tree_log), itw)
        got_c = interpolate(vals, L, itw, tree_log)
        assert (exp_c == got_c).all(), ("interp", L)
        h_big = orc.canonic_half_coset(n)
        exp_e = orc.evaluate(exp_c, n, h_big[0], (init, tree_log), tw)
        got_e = evaluate(exp_c, L, n, tw, tree_log)
        assert (exp_e == got_e).all(), ("eval", L, blow)
        print("ok", L, blow, extra)

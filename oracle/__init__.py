"""ctypes wrapper around oracle/tstwo_oracle.c -- the CPU restatement of TStwo's CpuBackend.

TEST INFRASTRUCTURE ONLY.  Importable from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  Nothing under tstwo_b200/ may import this package.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_DIR = os.path.dirname(os.path.abspath(__file__))
_SO = os.environ.get("TSTWO_ORACLE_SO") or os.path.join(_DIR, "_build", "libtstwo_oracle.so")   # scripts/oracle_asan.sh overrides

P = 2147483647
u32p = C.POINTER(C.c_uint32)
u8p = C.POINTER(C.c_uint8)


def build(force: bool = False) -> str:
    src = os.path.join(_DIR, "tstwo_oracle.c")
    if os.environ.get("TSTWO_ORACLE_SO"):
        return _SO
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _DIR, "-B" if force else "-s"], stdout=subprocess.DEVNULL)
    return _SO


def _cpu_stamp() -> str:
    import hashlib
    try:
        txt = open("/proc/cpuinfo").read()
        keep = [l for l in txt.splitlines() if l.startswith(("model name", "flags"))][:2]
    except OSError:
        keep = []
    return hashlib.sha1("\n".join(keep).encode()).hexdigest()[:16]


def use_native_build() -> bool:
    """bench.py's CPU legs: rebuild the SAME source with -O3 -march=native for the host it runs on (the shipped .so is
    -march=x86-64-v2 so that one build runs everywhere; a -march=native build made on another CPU may not).  Must be called
    before the first lib() use.  Returns False (and keeps the portable build) if the compiler is missing or fails."""
    global _SO, _lib
    if _lib is not None:
        return _SO.endswith("_native.so")
    so = os.path.join(_DIR, "_build", "libtstwo_oracle_native.so")
    stamp = so + ".cpu"
    src = os.path.join(_DIR, "tstwo_oracle.c")
    try:
        fresh = (os.path.exists(so) and os.path.exists(stamp) and open(stamp).read() == _cpu_stamp()
                 and os.path.getmtime(so) >= os.path.getmtime(src))
        if not fresh:
            os.makedirs(os.path.dirname(so), exist_ok=True)
            subprocess.check_call(["/usr/bin/gcc", "-O3", "-march=native", "-fPIC", "-fvisibility=hidden", "-fopenmp", "-shared",
                                   "-o", so, src], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            open(stamp, "w").write(_cpu_stamp())
        _SO = so
        return True
    except Exception:
        return False


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        L = _lib
        for name in ("orc_m31_add", "orc_m31_sub", "orc_m31_mul"):
            getattr(L, name).restype = C.c_uint32
            getattr(L, name).argtypes = [C.c_uint32, C.c_uint32]
        for name in ("orc_m31_neg", "orc_m31_inv", "orc_m31_partial_reduce"):
            getattr(L, name).restype = C.c_uint32
            getattr(L, name).argtypes = [C.c_uint32]
        L.orc_m31_reduce.restype = C.c_uint32
        L.orc_m31_reduce.argtypes = [C.c_uint64]
        L.orc_domain_index_at.restype = C.c_uint32
        L.orc_domain_index_at.argtypes = [C.c_uint32, C.c_int, C.c_uint32]
        L.orc_domain_iter_index.restype = C.c_uint32
        L.orc_domain_iter_index.argtypes = [C.c_uint32, C.c_int, C.c_uint32]
    return _lib


def _p(a: np.ndarray):
    assert a.dtype == np.uint32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(u32p)


def _u32(a) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a, dtype=np.uint32))


# ---- circle / coset index helpers (pure python; circle.ts:152-291, canonic.ts, domain.ts) ----
def subgroup_gen(log_size: int) -> int:
    return 0 if log_size == 0 else 1 << (31 - log_size)


def half_odds_index(log_size: int) -> int:
    """Coset.half_odds(log_size).initial_index (circle.ts:232-234)."""
    return subgroup_gen(log_size + 2)


def canonic_half_coset(log_size: int):
    """CanonicCoset(log_size).half_coset() -> (initial_index, log) (canonic.ts half_coset)."""
    return half_odds_index(log_size - 1), log_size - 1


def index_to_point(index: int):
    xy = np.zeros(2, dtype=np.uint32)
    lib().orc_index_to_point(C.c_uint32(index & 0x7FFFFFFF), _p(xy))
    return int(xy[0]), int(xy[1])


def domain_point(half_init: int, half_log: int, i: int, iter_order: bool = False):
    f = lib().orc_domain_iter_index if iter_order else lib().orc_domain_index_at
    return index_to_point(f(half_init, half_log, i))


def bit_reverse(a) -> np.ndarray:
    a = _u32(a).copy()
    if lib().orc_bit_reverse(_p(a), C.c_uint64(a.size)) != 0:
        raise ValueError("length is not power of two")
    return a


# ---- field vector ops ----
def cm31_op(op: int, a, b=None):
    a = _u32(a); out = np.zeros(2, dtype=np.uint32)
    lib().orc_cm31_op(op, _p(a), _p(_u32(b)) if b is not None else None, _p(out))
    return [int(v) for v in out]


def qm31_op(op: int, a, b=None):
    a = _u32(a); out = np.zeros(4, dtype=np.uint32)
    lib().orc_qm31_op(op, _p(a), _p(_u32(b)) if b is not None else None, _p(out))
    return [int(v) for v in out]


# ---- poly ops ----
def precompute_twiddles(initial_index: int, log_size: int):
    n = 1 << log_size
    tw = np.zeros(n, dtype=np.uint32); itw = np.zeros(n, dtype=np.uint32)
    lib().orc_precompute_twiddles(C.c_uint32(initial_index), log_size, _p(tw), _p(itw))
    return tw, itw


class TwiddleMismatch(Exception):
    pass


def evaluate(coeffs, domain_log: int, half_init: int, tw_root, tw, quirks: bool = True) -> np.ndarray:
    coeffs = _u32(coeffs)
    poly_log = coeffs.size.bit_length() - 1
    out = np.zeros(1 << domain_log, dtype=np.uint32)
    rc = lib().orc_evaluate(_p(coeffs), poly_log, _p(out), domain_log, C.c_uint32(half_init),
                            C.c_uint32(tw_root[0]), tw_root[1], _p(tw), int(quirks))
    if rc == -3:
        raise TwiddleMismatch("twiddle tree mismatch")
    if rc:
        raise ValueError(f"orc_evaluate rc={rc}")
    return out


def interpolate(values, half_init: int, tw_root, itw, quirks: bool = True) -> np.ndarray:
    v = _u32(values).copy()
    domain_log = v.size.bit_length() - 1
    rc = lib().orc_interpolate(_p(v), domain_log, C.c_uint32(half_init), C.c_uint32(tw_root[0]), tw_root[1],
                               _p(itw), int(quirks))
    if rc == -3:
        raise TwiddleMismatch("twiddle tree mismatch")
    if rc:
        raise ValueError(f"orc_interpolate rc={rc}")
    return v


def eval_at_point(coeffs, px, py):
    coeffs = _u32(coeffs)
    out = np.zeros(4, dtype=np.uint32)
    lib().orc_eval_at_point(_p(coeffs), coeffs.size.bit_length() - 1, _p(_u32(px)), _p(_u32(py)), _p(out))
    return [int(v) for v in out]


# ---- blake2s / merkle ----
def blake2s(msg: bytes) -> bytes:
    out = (C.c_uint8 * 32)()
    buf = (C.c_uint8 * max(1, len(msg))).from_buffer_copy(msg if msg else b"\0")
    lib().orc_blake2s(buf, C.c_uint64(len(msg)), out)
    return bytes(out)


def merkle_commit_layer(log_size: int, prev, cols) -> np.ndarray:
    cols = [_u32(c) for c in cols]
    arr = (u32p * max(1, len(cols)))(*[_p(c) for c in cols])
    out = np.zeros((1 << log_size, 32), dtype=np.uint8)
    pp = None
    if prev is not None:
        prev = np.ascontiguousarray(prev, dtype=np.uint8)
        assert prev.size == 64 << log_size
        pp = prev.ctypes.data_as(u8p)
    lib().orc_merkle_commit_layer(log_size, pp, arr, len(cols), out.ctypes.data_as(u8p))
    return out


def merkle_commit(cols):
    """MerkleProver.commit: returns list of layers, layers[0] = root layer (1 hash), as (2^k, 32) uint8."""
    cols = [_u32(c) for c in cols]
    if not cols:
        return [merkle_commit_layer(0, None, [])]
    logs = [c.size.bit_length() - 1 for c in cols]
    max_log = max(logs)
    arr = (u32p * len(cols))(*[_p(c) for c in cols])
    la = (C.c_int * len(cols))(*logs)
    out = np.zeros(((2 << max_log) - 1, 32), dtype=np.uint8)
    lib().orc_merkle_commit(arr, la, len(cols), out.ctypes.data_as(u8p))
    return [out[(1 << k) - 1:(2 << k) - 1] for k in range(max_log + 1)]


# ---- FRI ----
def fold_line(src4, coset_init: int, alpha):
    s = [_u32(c) for c in src4]
    log_size = s[0].size.bit_length() - 1
    d = [np.zeros(s[0].size // 2, dtype=np.uint32) for _ in range(4)]
    rc = lib().orc_fold_line(*[_p(c) for c in s], log_size, C.c_uint32(coset_init), _p(_u32(alpha)), *[_p(c) for c in d])
    if rc:
        raise ValueError("fold_line: Evaluation too small, must have at least 2 elements.")
    return d


def fold_circle_into_line(dst4, src4, half_init: int, alpha):
    s = [_u32(c) for c in src4]
    d = [_u32(c).copy() for c in dst4]
    log_size = s[0].size.bit_length() - 1
    if d[0].size * 2 != s[0].size:
        raise ValueError("fold_circle_into_line: Length mismatch between src and dst after considering fold step.")
    lib().orc_fold_circle_into_line(*[_p(c) for c in d], *[_p(c) for c in s], log_size, C.c_uint32(half_init), _p(_u32(alpha)))
    return d


def fri_decompose(src4):
    s = [_u32(c) for c in src4]
    log_size = s[0].size.bit_length() - 1
    g = [np.zeros_like(s[0]) for _ in range(4)]
    lam = np.zeros(4, dtype=np.uint32)
    lib().orc_fri_decompose(*[_p(c) for c in s], log_size, *[_p(c) for c in g], _p(lam))
    return g, [int(v) for v in lam]


# ---- accumulation ----
def accumulate(dst4, src4):
    d = [_u32(c).copy() for c in dst4]; s = [_u32(c) for c in src4]
    if d[0].size != s[0].size:
        raise ValueError("column length mismatch")
    lib().orc_accumulate(*[_p(c) for c in d], *[_p(c) for c in s], C.c_uint64(d[0].size))
    return d


def generate_secure_powers(felt, n: int):
    out = np.zeros(4 * max(n, 1), dtype=np.uint32)
    lib().orc_generate_secure_powers(_p(_u32(felt)), n, _p(out))
    return out[:4 * n].reshape(n, 4)


def accumulate_quotients(log_size: int, half_init: int, cols, random_coeff, batches, ts_denominators: bool = True):
    """batches: list of (point=(px[4], py[4]), [(col_index, value[4]), ...]).  ts_denominators=False: Stwo's CM31 denominators."""
    lib().orc_set_quotient_ts_denominators(1 if ts_denominators else 0)
    cols = [_u32(c) for c in cols]
    arr = (u32p * len(cols))(*[_p(c) for c in cols])
    pts = _u32([list(b[0][0]) + list(b[0][1]) for b in batches]).reshape(-1)
    offs = [0]
    ci, vals = [], []
    for b in batches:
        for (c, v) in b[1]:
            ci.append(c); vals.extend(v)
        offs.append(len(ci))
    offs_a = (C.c_int * len(offs))(*offs)
    ci_a = (C.c_int * max(1, len(ci)))(*ci)
    vals = _u32(vals if vals else [0])
    out = [np.zeros(1 << log_size, dtype=np.uint32) for _ in range(4)]
    lib().orc_accumulate_quotients(log_size, C.c_uint32(half_init), arr, len(cols), _p(_u32(random_coeff)), len(batches),
                                   _p(pts), offs_a, ci_a, _p(vals), *[_p(o) for o in out])
    return out


def coset_vanishing(coset_initial_index: int, coset_log_size: int, p):
    """constraints.ts:32-47 for an M31 point p = (x, y): x-coordinate of (p - initial + half_step), doubled log_size - 1 times."""
    P = 2147483647
    ix, iy = index_to_point(coset_initial_index)
    hx, hy = index_to_point(subgroup_gen(coset_log_size) >> 1)
    sx, sy = (p[0] * ix + p[1] * iy) % P, (p[1] * ix - p[0] * iy) % P      # p - initial
    x = (sx * hx - sy * hy) % P                                            # + half, x only
    for _ in range(1, coset_log_size):
        x = (2 * x * x - 1) % P
    return x


def widefib_denom_inv(trace_log: int, eval_log: int):
    """Stwo FrameworkComponent::evaluate_constraint_quotients_on_domain: inverse of the vanishing polynomial of
    CanonicCoset(trace_log).coset (Coset.odds: initial G_{2^(log+1)}, canonic.ts:24-40) at eval_domain.at(i), i < 2^(eval_log - trace_log),
    then bit-reversed."""
    P = 2147483647
    hi, hl = canonic_half_coset(eval_log)
    vals = [pow(coset_vanishing(subgroup_gen(trace_log + 1), trace_log, domain_point(hi, hl, i)), P - 2, P)
            for i in range(1 << (eval_log - trace_log))]
    return _bitrev_list(vals)


def widefib_constraints(trace_log: int, eval_log: int, cols, coeffs, denom_inv, out4):
    """out4 (4 uint32 arrays) += denom_inv[row >> trace_log] * sum_i coeffs[i] (c - a^2 - b^2)_i(row).  parity unpinned."""
    cols = [_u32(c) for c in cols]
    arr = (u32p * len(cols))(*[_p(c) for c in cols])
    co = _u32([x for q in coeffs for x in q] or [0])
    out = [_u32(o).copy() for o in out4]
    lib().orc_widefib_constraints(trace_log, eval_log, arr, len(cols), _p(co), _p(_u32(denom_inv)), *[_p(o) for o in out])
    return out


# ---- the metric pipeline ----
def lde_commit(cols_in: np.ndarray, log_blowup: int, tw, itw, want_coeffs=False, want_layers=True, quirks=True):
    """cols_in: (n_cols, 2^log) uint32.  Returns (coeffs|None, evals, layers|None)."""
    cols_in = _u32(cols_in)
    n_cols, n = cols_in.shape
    log_size = n.bit_length() - 1
    ext = log_size + log_blowup
    evals = np.zeros((n_cols, 1 << ext), dtype=np.uint32)
    coeffs = np.zeros((n_cols, n), dtype=np.uint32) if want_coeffs else None
    layers = np.zeros(((2 << ext) - 1, 32), dtype=np.uint8) if want_layers else None
    rc = lib().orc_lde_commit(_p(cols_in), n_cols, log_size, log_blowup, _p(tw), _p(itw),
                              _p(coeffs) if want_coeffs else None, _p(evals),
                              layers.ctypes.data_as(u8p) if want_layers else None, int(quirks))
    if rc:
        raise ValueError(f"orc_lde_commit rc={rc}")
    lay = [layers[(1 << k) - 1:(2 << k) - 1] for k in range(ext + 1)] if want_layers else None
    return coeffs, evals, lay


def num_threads() -> int:
    return int(lib().orc_num_threads())


def use_all_cores() -> int:
    """OpenMP threads = the cores this process may run on (torchrun sets OMP_NUM_THREADS=1 for its workers)."""
    import os
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    lib().orc_set_num_threads(int(n))
    return num_threads()


# ------------------------------------------------------------------------------------------------ Fiat-Shamir channel
class Channel:
    """CPU restatement of Blake2sChannel (core/src/channel/blake2.ts:25-224) and Blake2sMerkleChannel.mix_root
    (core/src/vcs/blake2_merkle.ts:28-31) over the oracle's own BLAKE2s (orc_blake2s).  Pinned by the two digests the
    reference's tests took from Stwo (core/test/channel/channel_exact_rust_tests.test.ts:97-118).  Test infrastructure only.

    State: 32-byte digest (zeros at creation, blake2_hash.ts), ChannelTime (n_challenges, n_sent; every digest update
    increments n_challenges and resets n_sent, channel/index.ts:97-104) and the queue of base-field draws that draw_felt
    keeps between calls (blake2.ts:172-181; draw_felts uses a fresh local queue, :183-205)."""

    P = 2147483647

    def __init__(self):
        self.digest = bytes(32)
        self.n_challenges = 0
        self.n_sent = 0
        self.base_queue = []

    def _update(self, new_digest: bytes):          # blake2.ts:75-79
        self.digest = new_digest
        self.n_challenges += 1
        self.n_sent = 0

    def mix_root(self, root: bytes):               # blake2_merkle.ts:29-31: H(digest || root)
        self._update(blake2s(self.digest + bytes(root)))

    def mix_felts(self, felts):                    # blake2.ts:113-118: H(digest || LE32 of every coordinate)
        words = [int(w) for f in felts for w in f]
        self._update(blake2s(self.digest + b"".join(int(w).to_bytes(4, "little") for w in words)))

    def mix_u32s(self, data):                      # blake2.ts:120-137
        self._update(blake2s(self.digest + b"".join(int(w).to_bytes(4, "little") for w in data)))

    def mix_u64(self, value: int):                 # blake2.ts:139-151: low word first
        self.mix_u32s([value & 0xFFFFFFFF, (value >> 32) & 0xFFFFFFFF])

    def draw_random_bytes(self) -> bytes:          # blake2.ts:207-221: H(digest || LE32(n_sent) || 28 zero bytes)
        out = blake2s(self.digest + self.n_sent.to_bytes(4, "little") + bytes(28))
        self.n_sent += 1
        return out

    def _draw_base_felts(self):                    # blake2.ts:157-170: 8 words, retry unless all < 2P, then reduce
        while True:
            b = self.draw_random_bytes()
            u = [int.from_bytes(b[4 * i:4 * i + 4], "little") for i in range(8)]
            if all(x < 2 * self.P for x in u):
                return [x % self.P for x in u]

    def draw_felt(self):                           # blake2.ts:172-181
        while len(self.base_queue) < 4:
            self.base_queue += self._draw_base_felts()
        out, self.base_queue = self.base_queue[:4], self.base_queue[4:]
        return out

    def draw_felts(self, n: int):                  # blake2.ts:183-205
        out, queue = [], []
        for _ in range(n):
            while len(queue) < 4:
                queue += self._draw_base_felts()
            out.append(queue[:4]); queue = queue[4:]
        return out

    def trailing_zeros(self) -> int:               # blake2.ts:94-111: of the low 128 bits, little endian
        v = int.from_bytes(self.digest[:16], "little")
        return 128 if v == 0 else (v & -v).bit_length() - 1


# ------------------------------------------------------------------------------------------------ Merkle decommit / verify
def _next_node(last, pi, direct, di):
    """nextDecommitmentNode, core/src/vcs/utils.ts:40-57."""
    cand = []
    if pi < len(last): cand.append(last[pi] // 2)
    if di < len(direct): cand.append(direct[di])
    return min(cand) if cand else None


def merkle_decommit(layers, queries, cols):
    """MerkleProver.decommit, core/src/vcs/prover.ts:32-109.  layers[k]: (2^k, 32) uint8 as returned by merkle_commit;
    queries: {log_size: sorted positions}; cols: the committed columns.  Returns (queried, hash_witness, column_witness)."""
    order = sorted(range(len(cols)), key=lambda i: -len(cols[i]))
    queried, hash_w, col_w = [], [], []
    ci, last = 0, []
    for log in range(len(layers) - 1, -1, -1):
        layer_cols = []
        while ci < len(order) and len(cols[order[ci]]) == 1 << log:
            layer_cols.append(cols[order[ci]]); ci += 1
        child = layers[log + 1] if log + 1 < len(layers) else None
        direct = list(queries.get(log, []))
        pi = di = 0
        nxt = []
        while (node := _next_node(last, pi, direct, di)) is not None:
            if child is not None:
                for ch in (2 * node, 2 * node + 1):
                    if pi < len(last) and last[pi] == ch: pi += 1
                    else: hash_w.append(bytes(child[ch]))
            vals = [int(c[node]) for c in layer_cols]
            if di < len(direct) and direct[di] == node:
                di += 1; queried += vals
            else:
                col_w += vals
            nxt.append(node)
        last = nxt
    return queried, hash_w, col_w


def merkle_verify(root: bytes, column_log_sizes, queries, queried, hash_w, col_w):
    """MerkleVerifier.verify, core/src/vcs/verifier.ts:28-150, with Blake2sMerkleHasher.hashNode
    (vcs/blake2_merkle.ts:9-24).  Raises ValueError with the reference's error names."""
    if not column_log_sizes:
        return
    n_cols = {}
    for l in column_log_sizes: n_cols[l] = n_cols.get(l, 0) + 1
    qi = hi = wi = 0
    last = None
    for log in range(max(column_log_sizes), -1, -1):
        prev_q = [q for q, _ in last] if last else []
        direct = list(queries.get(log, []))
        pi = di = 0       # pi walks prev_q for node selection, hp walks (index, hash) pairs for child look-up
        hp = 0
        cur = []
        while (node := _next_node(prev_q, pi, direct, di)) is not None:
            while pi < len(prev_q) and prev_q[pi] // 2 == node: pi += 1
            msg = b""
            if last is not None:
                for ch in (2 * node, 2 * node + 1):
                    if hp < len(last) and last[hp][0] == ch:
                        msg += last[hp][1]; hp += 1
                    else:
                        if hi >= len(hash_w): raise ValueError("WitnessTooShort")
                        msg += hash_w[hi]; hi += 1
            from_q = di < len(direct) and direct[di] == node
            if from_q: di += 1
            for _ in range(n_cols.get(log, 0)):
                if from_q:
                    if qi >= len(queried): raise ValueError("TooFewQueriedValues")
                    msg += int(queried[qi]).to_bytes(4, "little"); qi += 1
                else:
                    if wi >= len(col_w): raise ValueError("WitnessTooShort")
                    msg += int(col_w[wi]).to_bytes(4, "little"); wi += 1
            cur.append((node, blake2s(msg)))
        last = cur
    if hi != len(hash_w): raise ValueError("WitnessTooLong")
    if qi != len(queried): raise ValueError("TooManyQueriedValues")
    if wi != len(col_w): raise ValueError("WitnessTooLong")
    if last[0][1] != bytes(root): raise ValueError("RootMismatch")


# ------------------------------------------------------------------------------------------------ FRI commit / decommit flow
def _bitrev_list(items):
    n = len(items); log = n.bit_length() - 1
    return [items[int(format(i, f"0{log}b")[::-1], 2) if log else 0] for i in range(n)]


def line_interpolate(coset_init: int, log_size: int, coords):
    """LineEvaluation.interpolate, core/src/poly/line.ts:311-328 with lineIfft :354-390: bit-reverse the evaluations, butterfly
    layers with 1/x of the (repeatedly doubled) line domain, scale by 1/n.  Returns bit-reversed-order coefficients (n x 4)."""
    Pm = 2147483647
    n = 1 << log_size
    vals = [_bitrev_list([int(v) for v in c]) for c in coords]
    init, step, size = coset_init, subgroup_gen(log_size), n
    while size > 1:
        half = size // 2
        inv = [pow(index_to_point(init + i * step)[0], Pm - 2, Pm) for i in range(half)]
        for v in vals:
            for start in range(0, n, size):
                for i in range(half):
                    a, b = v[start + i], v[start + i + half]
                    v[start + i], v[start + i + half] = (a + b) % Pm, (a - b) * inv[i] % Pm
        init, step, size = (2 * init) & 0x7FFFFFFF, (2 * step) & 0x7FFFFFFF, half      # Coset.double, circle.ts:268-276
    n_inv = pow(n, Pm - 2, Pm)
    return [[vals[k][i] * n_inv % Pm for k in range(4)] for i in range(n)]


def fri_decommitment_positions(query_positions, fold_step):
    """computeDecommitmentPositionsAndWitnessEvals (core/src/fri.ts:346-384), positions only: (all positions, witness positions)."""
    positions, witness = [], []
    i = 0
    while i < len(query_positions):
        coset = query_positions[i] >> fold_step
        here = set()
        while i < len(query_positions) and query_positions[i] >> fold_step == coset:
            here.add(query_positions[i]); i += 1
        for pos in range(coset << fold_step, (coset + 1) << fold_step):
            positions.append(pos)
            if pos not in here: witness.append(pos)
    return positions, witness


def fri_commit(channel: "Channel", log_last_layer_degree_bound: int, log_blowup: int, columns):
    """FriProver.commit (core/src/fri.ts:645-754) with the placeholders filled in as in Stwo: first-layer tree over all coordinate
    columns, mix_root, alpha = draw_felt, fold_circle_into_line; per inner layer: commit, mix_root, draw alpha, fold_line, fold in
    the circle column whose folded size matches; last layer interpolated and mixed as felts.
    columns: [(log_size, [4 coordinate arrays])] on canonic domains, strictly decreasing.  Returns a dict of everything a test
    wants to compare."""
    coord_cols = [c for _, col in columns for c in col]
    first_layers = merkle_commit(coord_cols)
    channel.mix_root(first_layers[0][0].tobytes())
    alphas = [channel.draw_felt()]
    first_log = columns[0][0] - 1
    line_init, line_log = half_odds_index(first_log), first_log
    zero = [np.zeros(1 << first_log, dtype=np.uint32) for _ in range(4)]
    layer_eval = fold_circle_into_line(zero, columns[0][1], canonic_half_coset(columns[0][0])[0], alphas[0])
    pending = list(columns[1:])
    inner = []
    last_size = 1 << (log_last_layer_degree_bound + log_blowup)
    while len(layer_eval[0]) > last_size:
        trees = merkle_commit(layer_eval)
        channel.mix_root(trees[0][0].tobytes())
        alpha = channel.draw_felt(); alphas.append(alpha)
        inner.append({"evaluation": layer_eval, "layers": trees, "log": line_log})
        layer_eval = fold_line(layer_eval, line_init, alpha)
        line_init, line_log = (2 * line_init) & 0x7FFFFFFF, line_log - 1
        if pending and (1 << pending[0][0]) >> 1 == len(layer_eval[0]):
            log, col = pending.pop(0)
            layer_eval = fold_circle_into_line(layer_eval, col, canonic_half_coset(log)[0], alpha)
    assert not pending
    assert len(layer_eval[0]) == last_size, "last layer domain size mismatch"
    coeffs_br = line_interpolate(line_init, line_log, layer_eval)
    ordered = _bitrev_list(coeffs_br)
    bound = 1 << log_last_layer_degree_bound
    if any(any(c) for c in ordered[bound:]):
        raise ValueError("invalid degree")
    last_poly = _bitrev_list(ordered[:bound])
    channel.mix_felts(last_poly)
    return {"columns": columns, "first_layers": first_layers, "inner": inner, "alphas": alphas, "last_poly": last_poly,
            "last_eval": layer_eval}


def fri_decommit_on_queries(state, positions, log_domain_size):
    """FriProver.decommitOnQueries (core/src/fri.ts:769-785) with the first/inner layer decommit of Stwo in place of the TS
    placeholders (:509-541).  Returns [(fri_witness, (queried, hash_witness, column_witness), root)] first layer first."""
    fold = lambda pos, k: sorted({p >> k for p in pos})
    out = []
    by_log, witness = {}, []
    for log, col in state["columns"]:
        col_q = fold(positions, log_domain_size - log)
        pos, wit = fri_decommitment_positions(col_q, 1)
        by_log[log] = pos
        witness += [[int(c[p]) for c in col] for p in wit]
    coord_cols = [c for _, col in state["columns"] for c in col]
    out.append((witness, merkle_decommit(state["first_layers"], by_log, coord_cols), state["first_layers"][0][0].tobytes()))
    q = fold(positions, 1)
    for layer in state["inner"]:
        pos, wit = fri_decommitment_positions(q, 1)
        w = [[int(c[p]) for c in layer["evaluation"]] for p in wit]
        out.append((w, merkle_decommit(layer["layers"], {layer["log"]: pos}, layer["evaluation"]), layer["layers"][0][0].tobytes()))
        q = fold(q, 1)
    return out


def grind(channel: "Channel", pow_bits: int) -> int:
    """GrindOps.grind, core/src/backend/cpu/grind.ts:31-42: clone, mix_u64(nonce), trailing_zeros() >= powBits, else nonce += 1."""
    nonce = 0
    while True:
        c = Channel(); c.digest = channel.digest
        c.mix_u64(nonce)
        if c.trailing_zeros() >= pow_bits:
            return nonce
        nonce += 1

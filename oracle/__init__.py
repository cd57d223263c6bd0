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
_SO = os.path.join(_DIR, "_build", "libtstwo_oracle.so")

P = 2147483647
u32p = C.POINTER(C.c_uint32)
u8p = C.POINTER(C.c_uint8)


def build(force: bool = False) -> str:
    src = os.path.join(_DIR, "tstwo_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _DIR, "-B" if force else "-s"], stdout=subprocess.DEVNULL)
    return _SO


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


def accumulate_quotients(log_size: int, half_init: int, cols, random_coeff, batches):
    """batches: list of (point=(px[4], py[4]), [(col_index, value[4]), ...])."""
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

"""Parity at the sizes BASELINE.json names, against the ORACLE (never against another CUDA path), and the reference's golden
field vectors replayed on the DEVICE arithmetic.

  * test_field_vectors_on_device   test-vectors/{m31,cm31,qm31}-test-vectors.json (tests/golden/field_vectors.json) through
                                   tstwo_debug_field_op, i.e. csrc/m31.cuh one operation per thread -- the pattern of
                                   test-equivalence/fields/{m31,cm31,qm31}.test.ts
  * config 1 verbatim              1 column, log_size 16, blow-up 4, LDE + Blake2s Merkle commit, every array in full
  * config 3                       256 x 2^22, blow-up 2 on bench.py's own trace: EVERY evaluation column and the root
                                   against the oracle; pins the root bench.py prints
  * config 4                       fold_circle_into_line on a 2^24 QM31 column, then every fold_line down to 2 elements
                                   (fri.ts:120-192, 676-716), each layer in full, with and without a twiddle tree
"""
import ctypes as C
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
P = 2**31 - 1
G = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "field_vectors.json")))


@pytest.fixture(scope="module")
def tb():
    import tstwo_b200
    return tstwo_b200


def _field_op(L, chk, field, op, a_rows, b_rows=None):
    n = len(a_rows)
    a = np.zeros((n, 4), dtype=np.uint32)
    b = np.zeros((n, 4), dtype=np.uint32)
    for i, r in enumerate(a_rows):
        a[i, :len(r)] = r
    if b_rows is not None:
        for i, r in enumerate(b_rows):
            b[i, :len(r)] = r
    out = np.zeros((n, 4), dtype=np.uint32)
    u32p = C.POINTER(C.c_uint32)
    chk(L.tstwo_debug_field_op(field, op, a.ctypes.data_as(u32p), b.ctypes.data_as(u32p), out.ctypes.data_as(u32p), n))
    return out


def test_field_vectors_on_device(tb):
    from tstwo_b200 import _lib
    L = _lib.init(0)
    chk = _lib.check
    # ---- M31 (452 vectors)
    n = 0
    by = {}
    for v in G["m31"]:
        by.setdefault(v["op"], []).append(v)
    for name, op in (("add", 0), ("sub", 1), ("mul", 2)):
        vs = by[name]
        got = _field_op(L, chk, 0, op, [[v["in"]["a"]] for v in vs], [[v["in"]["b"]] for v in vs])
        assert got[:, 0].tolist() == [v["out"] for v in vs], name
        n += len(vs)
    vs = by["neg"]
    assert _field_op(L, chk, 0, 3, [[v["in"]["a"]] for v in vs])[:, 0].tolist() == [v["out"] for v in vs]
    n += len(vs)
    for name in ("inverse", "pow2147483645"):
        vs = by.get(name, [])
        if vs:
            assert _field_op(L, chk, 0, 4, [[v["in"]["value"]] for v in vs])[:, 0].tolist() == [v["out"] for v in vs], name
            n += len(vs)
    for name in ("reduce", "from_u32"):
        vs = by.get(name, [])
        if vs:
            vals = [int(v["in"]["value"]) for v in vs]
            got = _field_op(L, chk, 0, 7, [[x & 0xFFFFFFFF, x >> 32] for x in vals])
            assert got[:, 0].tolist() == [v["out"] for v in vs], name
            n += len(vs)
    vs = by.get("from_i32", [])
    if vs:   # m31.ts `from`: negative values are 2P - |v| before the reduction
        vals = [int(v["in"]["value"]) for v in vs]
        vals = [x if x >= 0 else 2 * P - abs(x) for x in vals]
        assert _field_op(L, chk, 0, 7, [[x & 0xFFFFFFFF, x >> 32] for x in vals])[:, 0].tolist() == [v["out"] for v in vs]
        n += len(vs)
    vs = by.get("partial_reduce", [])
    if vs:
        assert _field_op(L, chk, 0, 8, [[int(v["in"]["value"])] for v in vs])[:, 0].tolist() == [v["out"] for v in vs]
        n += len(vs)
    assert n >= 430, n
    # ---- CM31 (220 vectors)
    n = 0
    by = {}
    for v in G["cm31"]:
        by.setdefault(v["op"], []).append(v)
    for name, op in (("add", 0), ("sub", 1), ("mul", 2)):
        vs = by[name]
        got = _field_op(L, chk, 1, op, [[v["in"]["a_real"], v["in"]["a_imag"]] for v in vs], [[v["in"]["b_real"], v["in"]["b_imag"]] for v in vs])
        assert got[:, :2].tolist() == [[v["out"]["real"], v["out"]["imag"]] for v in vs], name
        n += len(vs)
    for name, op in (("neg", 3), ("inverse", 4)):
        vs = by[name]
        got = _field_op(L, chk, 1, op, [[v["in"]["real"], v["in"]["imag"]] for v in vs])
        assert got[:, :2].tolist() == [[v["out"]["real"], v["out"]["imag"]] for v in vs], name
        n += len(vs)
    assert n >= 205, n
    # ---- QM31 (135 vectors)
    n = 0
    by = {}
    for v in G["qm31"]:
        by.setdefault(v["op"], []).append(v)
    for name, op in (("add", 0), ("sub", 1), ("mul", 2)):
        vs = by[name]
        got = _field_op(L, chk, 2, op, [v["in"]["a"] for v in vs], [v["in"]["b"] for v in vs])
        assert got.tolist() == [v["out"] for v in vs], name
        n += len(vs)
    for name, op in (("neg", 3), ("inverse", 4)):
        vs = by[name]
        assert _field_op(L, chk, 2, op, [v["in"]["value"] for v in vs]).tolist() == [v["out"] for v in vs], name
        n += len(vs)
    vs = by.get("mul_cm31", [])
    if vs:
        got = _field_op(L, chk, 2, 5, [v["in"]["qm31"] for v in vs], [v["in"]["cm31"] for v in vs])
        assert got.tolist() == [v["out"] for v in vs]
        n += len(vs)
    vs = by.get("from_partial_evals", [])
    for v in vs:   # e0 + e1 * i + e2 * u + e3 * iu  (fields/qm31.ts from_partial_evals)
        e = v["in"]["evals"]
        r = np.array(e[0], dtype=np.uint32)
        for k, basis in enumerate([[0, 1, 0, 0], [0, 0, 1, 0], [0, 0, 0, 1]]):
            t = _field_op(L, chk, 2, 2, [e[k + 1]], [basis])[0]
            r = _field_op(L, chk, 2, 0, [r.tolist()], [t.tolist()])[0]
        assert r.tolist() == v["out"]
        n += 1
    assert n >= 125, n


def test_field_ops_random_against_oracle(tb, orc):
    """20 000 random operand pairs per operation against the oracle (itself pinned by the same golden vectors), incl. the
    edge residues 0, 1, P - 1."""
    from tstwo_b200 import _lib
    L = _lib.init(0)
    chk = _lib.check
    rng = np.random.default_rng(99)
    n = 20000
    a = rng.integers(0, P, (n, 4), dtype=np.uint32)
    b = rng.integers(0, P, (n, 4), dtype=np.uint32)
    a[:8] = [[0, 0, 0, 0], [1, 0, 0, 0], [P - 1] * 4, [0, 1, 0, 0], [0, 0, 1, 0], [P - 1, 0, 0, 0], [1, 1, 1, 1], [0, 0, 0, P - 1]]
    b[:8] = [[P - 1] * 4, [1, 0, 0, 0], [P - 1] * 4, [0, 1, 0, 0], [0, 0, 0, 1], [P - 1, 0, 0, 0], [0, 0, 0, 0], [2, 0, 0, 0]]
    lib = orc.lib()
    for op in (0, 1, 2):
        got = _field_op(L, chk, 2, op, a, b)
        for i in list(range(16)) + list(rng.integers(0, n, 300)):
            assert got[i].tolist() == orc.qm31_op(op, a[i], b[i]), (op, i)
        gm = _field_op(L, chk, 0, op, a[:, :1], b[:, :1])[:, 0]
        f = (lib.orc_m31_add, lib.orc_m31_sub, lib.orc_m31_mul)[op]
        for i in list(range(16)) + list(rng.integers(0, n, 300)):
            assert int(gm[i]) == f(int(a[i, 0]), int(b[i, 0])), (op, i)
    nz = a.copy()
    nz[(nz == 0).all(axis=1)] = 1
    got = _field_op(L, chk, 2, 4, nz)
    for i in list(range(16)) + list(rng.integers(0, n, 200)):
        assert got[i].tolist() == orc.qm31_op(4, nz[i]), i
    # inverse times the element is one, for all 20 000
    prod = _field_op(L, chk, 2, 2, nz, got)
    assert (prod == np.array([1, 0, 0, 0], dtype=np.uint32)).all()


def _splitmix64(seed, n):
    out = np.empty(n, dtype=np.uint64)
    x = np.uint64(seed)
    with np.errstate(over="ignore"):
        for i in range(n):
            x = x + np.uint64(0x9E3779B97F4A7C15)
            z = x
            z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
            z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
            out[i] = z ^ (z >> np.uint64(31))
    return out


def test_config1_verbatim(tb, orc):
    """BASELINE config 1 / SURVEY 8(d) C1: one column, log_size 16, blow-up 4, x_i = splitmix64(0x5eed0001, i) mod P taken as
    evaluations in bit-reversed order on CanonicCoset(16).circleDomain(); coefficients, evaluations, EVERY Merkle layer."""
    log, blow = 16, 2
    ext = log + blow
    vals = (_splitmix64(0x5EED0001, 1 << log) % np.uint64(P)).astype(np.uint32).reshape(1, -1)
    root_coset = tb.CanonicCoset.new(ext).half_coset()
    tw = tb.CudaBackend.precomputeTwiddles(root_coset)
    otw, oitw = orc.precompute_twiddles(root_coset.initial_index.value, root_coset.log_size)
    coeffs, evals, layers = orc.lde_commit(vals, blow, otw, oitw, want_coeffs=True)
    # (a) the fused entry point
    tree = tb.CommitmentTreeProver.from_trace(tb.CudaColumn.from_cpu(vals.reshape(-1)), 1, log, blow, tw)
    assert (tree.polynomials[0].coeffs.toCpu() == coeffs[0]).all()
    assert (tree.evaluations[0].values.toCpu() == evals[0]).all()
    for k in range(ext + 1):
        assert tree.commitment.layers[k].toCpu().tobytes() == layers[k].tobytes(), f"Merkle layer {k}"
    assert tree.root() == layers[0][0].tobytes()
    # (b) the reference's own call sequence: interpolate -> evaluate on the blow-up domain -> MerkleProver.commit
    dom = tb.CanonicCoset.new(log).circleDomain()
    poly = tb.CudaCircleEvaluation.new(dom, vals[0]).interpolateWithTwiddles(tw)
    assert (poly.coeffs.toCpu() == coeffs[0]).all()
    ev = poly.evaluateWithTwiddles(tb.CanonicCoset.new(ext).circleDomain(), tw)
    assert (ev.values.toCpu() == evals[0]).all()
    mp = tb.MerkleProver.commit(None, [ev.values])
    assert mp.root() == layers[0][0].tobytes()


def test_config3_every_column_and_root_against_oracle(tb, orc):
    """BASELINE config 3 / the bench.py job at FULL size: 256 columns of 2^22, blow-up 2, on the exact trace bench.py
    generates (bench.make_trace).  Every one of the 256 evaluation columns and the Merkle root are compared with the oracle
    (SURVEY 8d), and the root must be the one bench.py prints (bench.BENCH_ROOT) -- so the number in the driver's record is
    the throughput of a computation whose result is the oracle's."""
    import psutil
    import torch
    import bench
    from tstwo_b200 import _lib
    if psutil.virtual_memory().available < 26 * 2**30:
        pytest.skip("needs ~22 GiB of host memory for the oracle's copy of the job")
    L = _lib.init(0)
    log, blow, n_cols = bench.LOG_SIZE, bench.LOG_BLOWUP, bench.N_COLS
    n, ext = 1 << log, log + blow
    dev = torch.device("cuda", 0)
    trace = bench.make_trace(torch, dev, 0, 1, n_cols, n)
    host_trace = trace.cpu().numpy().view(np.uint32)
    assert int(host_trace.max()) < P
    root_coset = tb.CanonicCoset.new(ext).half_coset()
    tw = tb.CudaBackend.precomputeTwiddles(root_coset)
    coeffs = torch.empty((n_cols, n), dtype=torch.int32, device=dev)
    evals = torch.empty((n_cols, 1 << ext), dtype=torch.int32, device=dev)
    layers = torch.empty(((2 << ext) - 1) * 8, dtype=torch.int32, device=dev)
    root = (C.c_uint8 * 32)()
    torch.cuda.synchronize()
    _lib.check(L.tstwo_lde_commit(trace.data_ptr(), n_cols, log, blow, tw.handle, coeffs.data_ptr(), evals.data_ptr(),
                                  layers.data_ptr(), root, 0))
    _lib.check(L.tstwo_sync())
    assert bytes(root).hex() == bench.BENCH_ROOT, "the root of bench.py's job changed"
    del trace
    orc.use_all_cores()
    otw, oitw = orc.precompute_twiddles(root_coset.initial_index.value, root_coset.log_size)
    ocoeffs, oevals, olayers = orc.lde_commit(host_trace, blow, otw, oitw, want_coeffs=True)
    assert olayers[0][0].tobytes() == bytes(root), "Merkle root differs from the oracle's"
    for c in range(n_cols):
        assert (evals[c].cpu().numpy().view(np.uint32) == oevals[c]).all(), f"evaluation column {c}"
        assert (coeffs[c].cpu().numpy().view(np.uint32) == ocoeffs[c]).all(), f"coefficient column {c}"
    # and the leaf layer + the layer below it in full (the other layers follow from them and the root)
    got_layers = layers.cpu().numpy().view(np.uint8).reshape(-1, 32)
    for k in (ext, ext - 1, 3, 0):
        assert (got_layers[(1 << k) - 1:(2 << k) - 1] == olayers[k]).all(), f"Merkle layer {k}"


@pytest.mark.parametrize("with_tree", [True, False])
def test_config4_fri_fold_chain_full_size(tb, orc, with_tree):
    """BASELINE config 4 / SURVEY 8(d) C4: one QM31 column on CanonicCoset(24).circleDomain() (uniform random, seed
    0x5eed0004), alphas = 4 uniform M31 each, fold_circle_into_line then fold_line down to 2 elements (FriConfig(0, 1) ends
    there): every layer compared in full with the oracle.  with_tree = twiddles from the tree of the domain's half coset (the
    prover's path); without = generated on the fly (launch_coset_points at 2^23 entries)."""
    log = 24
    rng = np.random.default_rng(0x5EED0004)
    dom = tb.CanonicCoset.new(log).circleDomain()
    src = [rng.integers(0, P, 1 << log, dtype=np.uint32) for _ in range(4)]
    tw = tb.CudaBackend.precomputeTwiddles(dom.halfCoset) if with_tree else None
    orc.use_all_cores()
    alpha = [int(v) for v in rng.integers(1, P, 4)]
    layer = tb.LineEvaluation.newZero(tb.LineDomain.new(dom.halfCoset))
    tb.fold_circle_into_line(layer, tb.SecureEvaluation(dom, tb.SecureColumnByCoords.from_cpu(src)), alpha, tw)
    cur = orc.fold_circle_into_line([np.zeros(1 << (log - 1), dtype=np.uint32)] * 4, src, dom.halfCoset.initial_index.value, alpha)
    for k, (a, b) in enumerate(zip(layer.values.to_cpu(), cur)):
        assert (a == b).all(), f"circle fold, coordinate {k}"
    coset = dom.halfCoset
    n_layers = 0
    while layer.len() > 2:
        alpha = [int(v) for v in rng.integers(1, P, 4)]
        layer = tb.fold_line(layer, alpha, tw)
        cur = orc.fold_line(cur, coset.initial_index.value, alpha)
        coset = coset.double()
        n_layers += 1
        assert layer.len() == cur[0].size
        for k, (a, b) in enumerate(zip(layer.values.to_cpu(), cur)):
            assert (a == b).all(), f"line fold to 2^{coset.log_size}, coordinate {k}"
    assert n_layers == 22 and layer.len() == 2

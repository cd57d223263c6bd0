"""The metric path (tstwo_lde_commit / CommitmentTreeProver.from_trace) against the oracle at sizes the oracle
finishes in seconds, and through size-independent properties at BASELINE.json's config-2 size."""
import ctypes as C
import hashlib
import struct

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
P = 2147483647


@pytest.fixture(scope="module")
def tb():
    import tstwo_b200
    return tstwo_b200


@pytest.mark.parametrize("log,blow,n_cols,keep", [(4, 1, 3, True), (10, 2, 5, True), (14, 1, 24, True), (14, 1, 9, False),
                                                 (15, 2, 16, False), (16, 1, 64, True)])
def test_lde_commit_vs_oracle(tb, orc, log, blow, n_cols, keep):
    rng = np.random.default_rng(log * 100 + n_cols)
    trace = rng.integers(0, P, (n_cols, 1 << log), dtype=np.uint32)
    root_coset = tb.CanonicCoset.new(log + blow).half_coset()
    tw = tb.CudaBackend.precomputeTwiddles(root_coset)
    tree = tb.CommitmentTreeProver.from_trace(tb.CudaColumn.from_cpu(trace.reshape(-1)), n_cols, log, blow, tw, keep_coeffs=keep)
    otw, oitw = orc.precompute_twiddles(root_coset.initial_index.value, root_coset.log_size)
    coeffs, evals, layers = orc.lde_commit(trace, blow, otw, oitw, want_coeffs=True)
    for c in range(n_cols):
        if keep:
            assert (tree.polynomials[c].coeffs.toCpu() == coeffs[c]).all()
        assert (tree.evaluations[c].values.toCpu() == evals[c]).all()
    assert tree.root() == layers[0][0].tobytes()
    for k in (0, 1, log + blow - 1, log + blow):
        assert (tree.commitment.layer_bytes(k) == layers[k]).all()


@pytest.mark.parametrize("n_cols", [40, 16, 7, 33, 64])
def test_lde_commit_host_matches_device(tb, orc, n_cols):
    """Host-fed pipeline: the leaf hashes absorb 16 columns per PCIe chunk (ragged last chunk, fewer than one chunk,
    whole chunks) and must give the root of the device-resident path."""
    from tstwo_b200 import _lib
    L = _lib.init(0)
    log, blow = 14, 1
    rng = np.random.default_rng(5)
    trace = rng.integers(0, P, (n_cols, 1 << log), dtype=np.uint32)
    tw = tb.CudaBackend.precomputeTwiddles(tb.CanonicCoset.new(log + blow).half_coset())
    ref = tb.CommitmentTreeProver.from_trace(tb.CudaColumn.from_cpu(trace.reshape(-1)), n_cols, log, blow, tw)
    stage = tb.CudaColumn.uninitialized(trace.size)
    evals = tb.CudaColumn.uninitialized(trace.size << blow)
    layers = tb.CudaColumn.uninitialized(8 * ((2 << (log + blow)) - 1))
    root = (C.c_uint8 * 32)()
    _lib.check(L.tstwo_lde_commit_host(trace.ctypes.data, n_cols, log, blow, tw.handle, stage.ptr, 0, evals.ptr, layers.ptr, root, 1))
    assert bytes(root) == ref.root()


def test_config2_properties(tb, orc):
    """BASELINE config 2: wide-Fibonacci 256 columns, log 20, blowup 2.  Too big for an element-wise oracle compare in
    the test budget, so: (a) the first and last column are compared with the oracle in full, (b) a Merkle
    authentication path of a random leaf is re-hashed with hashlib up to the root, (c) interpolating the extended
    evaluation gives back the coefficients padded with zeros (round trip), (d) root is reproducible."""
    log, blow, n_cols = 20, 1, 256
    n = 1 << log
    # wide Fibonacci (examples/fibonacci.ts:245-258 a=1, b=row; c = a^2 + b^2), natural row order
    a = np.ones(n, dtype=np.uint64); b = np.arange(n, dtype=np.uint64) % P
    cols = np.empty((n_cols, n), dtype=np.uint32)
    cols[0], cols[1] = a, b
    for k in range(2, n_cols):
        a, b = b, (a * a % P + b * b % P) % P
        cols[k] = b
    root_coset = tb.CanonicCoset.new(log + blow).half_coset()
    tw = tb.CudaBackend.precomputeTwiddles(root_coset)
    tree = tb.CommitmentTreeProver.from_trace(tb.CudaColumn.from_cpu(cols.reshape(-1)), n_cols, log, blow, tw)
    otw, oitw = orc.precompute_twiddles(root_coset.initial_index.value, root_coset.log_size)
    for c in (0, n_cols - 1):
        cf, ev, _ = orc.lde_commit(cols[c:c + 1], blow, otw, oitw, want_coeffs=True, want_layers=False)
        assert (tree.polynomials[c].coeffs.toCpu() == cf[0]).all()
        assert (tree.evaluations[c].values.toCpu() == ev[0]).all()
    # (b) authentication path
    ext = log + blow
    leaf = 0x12345 % (1 << ext)
    row = b"".join(struct.pack("<I", tree.evaluations[c].values.at(leaf)) for c in range(n_cols))
    h = hashlib.blake2s(row).digest()
    idx = leaf
    for k in range(ext, 0, -1):
        layer = tree.commitment.layers[k]
        assert layer.slice(8 * idx, 8).toCpu().tobytes() == h
        sib = layer.slice(8 * (idx ^ 1), 8).toCpu().tobytes()
        h = hashlib.blake2s(h + sib if idx % 2 == 0 else sib + h).digest()
        idx >>= 1
    assert h == tree.root()
    # (c) round trip on the big domain
    back = tb.CudaBackend.interpolateColumns([tree.evaluations[7]], tw)[0].coeffs.toCpu()
    assert (back[:n] == tree.polynomials[7].coeffs.toCpu()).all() and not back[n:].any()
    # (d) determinism
    again = tb.CommitmentTreeProver.from_trace(tb.CudaColumn.from_cpu(cols.reshape(-1)), n_cols, log, blow, tw, keep_coeffs=False)
    assert again.root() == tree.root()


@pytest.mark.parametrize("n_shards", [2, 4, 8])
def test_lde_columns_sharded_layout(tb, orc, n_shards):
    """Multi-GPU building block: evaluations grouped by destination rank equal the plain LDE, re-laid out; hashing a
    row shard gives the matching subtree of the full Merkle tree (top layers = hash of the shard roots)."""
    from tstwo_b200 import _lib
    L = _lib.init(0)
    log, blow, n_cols = 14, 1, 12
    ext = log + blow
    rng = np.random.default_rng(n_shards)
    trace = rng.integers(0, P, (n_cols, 1 << log), dtype=np.uint32)
    tw = tb.CudaBackend.precomputeTwiddles(tb.CanonicCoset.new(ext).half_coset())
    ref = tb.CommitmentTreeProver.from_trace(tb.CudaColumn.from_cpu(trace.reshape(-1)), n_cols, log, blow, tw)
    full = np.stack([e.values.toCpu() for e in ref.evaluations])                     # (n_cols, 2^ext)
    dtrace = tb.CudaColumn.from_cpu(trace.reshape(-1))
    out = tb.CudaColumn.uninitialized(n_cols << ext)
    _lib.check(L.tstwo_lde_columns_sharded(dtrace.ptr, n_cols, log, blow, tw.handle, 0, out.ptr, n_shards, 0))
    rows = (1 << ext) // n_shards
    got = out.toCpu().reshape(n_shards, n_cols, rows)
    for s in range(n_shards):
        assert (got[s] == full[:, s * rows:(s + 1) * rows]).all()
    # subtree of shard 1 == the corresponding nodes of the full tree
    sub = ext - (n_shards.bit_length() - 1)
    shard = tb.CudaColumn.from_cpu(got[1].reshape(-1))
    mp = tb.MerkleProver.commit(tb.CudaBackend.merkleOps, [shard.slice(c * rows, rows) for c in range(n_cols)])
    k = n_shards.bit_length() - 1
    assert mp.root() == ref.commitment.layer_bytes(k)[1].tobytes()


@pytest.mark.parametrize("n_shards,col_offset,total", [(2, 0, 3), (4, 2, 6), (1, 1, 4)])
def test_lde_columns_to_peers_local_bases(tb, orc, n_shards, col_offset, total):
    """tstwo_lde_columns_to_peers with every shard base pointing at a LOCAL buffer (the multi-GPU run maps peers' buffers
    there): shard d must receive rows [d * rows, (d+1) * rows) of every evaluation at column col_offset + c, and nothing
    else may be touched."""
    from tstwo_b200 import _lib
    L = _lib.init(0)
    log, blow, n_cols = 16, 1, 3
    ext = log + blow
    rows = (1 << ext) // n_shards
    rng = np.random.default_rng(11 * n_shards + col_offset)
    trace = rng.integers(0, P, (n_cols, 1 << log), dtype=np.uint32)
    tw = tb.CudaBackend.precomputeTwiddles(tb.CanonicCoset.new(ext).half_coset())
    ref = tb.CommitmentTreeProver.from_trace(tb.CudaColumn.from_cpu(trace.reshape(-1)), n_cols, log, blow, tw)
    shards = [tb.CudaColumn.from_cpu(np.full(total * rows, 7, dtype=np.uint32)) for _ in range(n_shards)]
    work = tb.CudaColumn.uninitialized(n_cols << ext)
    _lib.check(L.tstwo_lde_columns_to_peers(tb.CudaColumn.from_cpu(trace.reshape(-1)).ptr, n_cols, log, blow, tw.handle, 0, work.ptr,
                                            _lib.dptr_array([s.ptr for s in shards]), n_shards, col_offset, 0))
    for d in range(n_shards):
        got = shards[d].toCpu().reshape(total, rows)
        for c in range(total):
            if col_offset <= c < col_offset + n_cols:
                want = ref.evaluations[c - col_offset].values.toCpu()[d * rows:(d + 1) * rows]
                assert (got[c] == want).all(), f"shard {d} column {c}"
            else:
                assert (got[c] == 7).all(), f"shard {d} column {c} must be untouched"


@pytest.mark.parametrize("groups,peer_sms,n_cols,taper", [(2, 48, 5, 100), (4, 60, 9, 100), (3, 100, 3, 100), (8, 16, 4, 100),
                                                          (4, 62, 16, 33), (5, 62, 13, 50), (4, 62, 5, 25), (3, 62, 3, 10)])
def test_lde_columns_to_peers_overlapped_matches(tb, orc, groups, peer_sms, n_cols, taper):
    """tstwo_lde_columns_to_peers_overlapped (last pass of each column group on a side stream with an SM budget) writes exactly
    what the one-group call writes, for uneven groups too (taper < 100: smaller first / last group, per-phase SM split); checked
    against the oracle's evaluations on local shard buffers."""
    from tstwo_b200 import _lib
    L = _lib.init(0)
    _lib.check(L.tstwo_debug_set_overlap_taper(taper))
    log, blow, n_shards, col_offset = 15, 1, 4, 1
    ext = log + blow
    rows = (1 << ext) // n_shards
    total = n_cols + 2
    rng = np.random.default_rng(100 * groups + n_cols)
    trace = rng.integers(0, P, (n_cols, 1 << log), dtype=np.uint32)
    tw = tb.CudaBackend.precomputeTwiddles(tb.CanonicCoset.new(ext).half_coset())
    root_coset = tb.CanonicCoset.new(ext).half_coset()
    otw, oitw = orc.precompute_twiddles(root_coset.initial_index.value, root_coset.log_size)
    _, evals, _ = orc.lde_commit(trace, blow, otw, oitw, want_coeffs=True)
    dtrace = tb.CudaColumn.from_cpu(trace.reshape(-1))
    for rep in range(3):                      # repeated calls reuse the side stream and its events
        shards = [tb.CudaColumn.from_cpu(np.full(total * rows, 7, dtype=np.uint32)) for _ in range(n_shards)]
        work = tb.CudaColumn.uninitialized(n_cols << ext)
        _lib.check(L.tstwo_lde_columns_to_peers_overlapped(dtrace.ptr, n_cols, log, blow, tw.handle, 0, work.ptr,
                                                           _lib.dptr_array([s.ptr for s in shards]), n_shards, col_offset, 0,
                                                           groups, peer_sms))
        for d in range(n_shards):
            got = shards[d].toCpu().reshape(total, rows)
            for c in range(total):
                if col_offset <= c < col_offset + n_cols:
                    assert (got[c] == evals[c - col_offset][d * rows:(d + 1) * rows]).all(), f"rep {rep} shard {d} column {c}"
                else:
                    assert (got[c] == 7).all(), f"shard {d} column {c} must be untouched"
    _lib.check(L.tstwo_debug_set_overlap_taper(100))


def test_lde_columns_to_peers_overlapped_rejects_bad_budget(tb):
    from tstwo_b200 import _lib
    L = _lib.init(0)
    tw = tb.CudaBackend.precomputeTwiddles(tb.CanonicCoset.new(16).half_coset())
    buf = tb.CudaColumn.uninitialized(1 << 17)
    rc = L.tstwo_lde_columns_to_peers_overlapped(buf.ptr, 1, 15, 1, tw.handle, 0, buf.ptr, _lib.dptr_array([buf.ptr]), 1, 0, 0, 2, 100000)
    assert rc != 0 and b"overlap" in L.tstwo_last_error()


def test_config3_full_size_single_and_sharded(tb, orc):
    """BASELINE config 3 at its FULL size on one GPU: 256 uniform-random columns of 2^22, blow-up 2.  (a) first / last
    column against the oracle in full, (b) an authentication path re-hashed with hashlib, (c) round trip, and (d) the
    multi-GPU data path emulated on one device: tstwo_lde_columns_to_peers_overlapped stores the rows of 8 shards into 8
    LOCAL leaf buffers (the multi-GPU run maps the peers' buffers there), each shard's subtree is committed on its own and the
    top log2(8) layers are finished as rank 0 does -- the root must equal the single-tree root."""
    from tstwo_b200 import _lib
    L = _lib.init(0)
    log, blow, n_cols, G = 22, 1, 256, 8
    n, ext = 1 << log, log + blow
    rng = np.random.default_rng(0x5EED0003)
    cols = rng.integers(0, P, (n_cols, n), dtype=np.uint32)
    root_coset = tb.CanonicCoset.new(ext).half_coset()
    tw = tb.CudaBackend.precomputeTwiddles(root_coset)
    dtrace = tb.CudaColumn.from_cpu(cols.reshape(-1))
    tree = tb.CommitmentTreeProver.from_trace(dtrace, n_cols, log, blow, tw)
    otw, oitw = orc.precompute_twiddles(root_coset.initial_index.value, root_coset.log_size)
    for c in (0, n_cols - 1):
        cf, ev, _ = orc.lde_commit(cols[c:c + 1], blow, otw, oitw, want_coeffs=True, want_layers=False)
        assert (tree.polynomials[c].coeffs.toCpu() == cf[0]).all()
        assert (tree.evaluations[c].values.toCpu() == ev[0]).all()
    leaf = 0x6543210 % (1 << ext)
    row = b"".join(struct.pack("<I", tree.evaluations[c].values.at(leaf)) for c in range(n_cols))
    h = hashlib.blake2s(row).digest()
    idx = leaf
    for k in range(ext, 0, -1):
        layer = tree.commitment.layers[k]
        assert layer.slice(8 * idx, 8).toCpu().tobytes() == h
        sib = layer.slice(8 * (idx ^ 1), 8).toCpu().tobytes()
        h = hashlib.blake2s(h + sib if idx % 2 == 0 else sib + h).digest()
        idx >>= 1
    assert h == tree.root()
    back = tb.CudaBackend.interpolateColumns([tree.evaluations[131]], tw)[0].coeffs.toCpu()
    assert (back[:n] == tree.polynomials[131].coeffs.toCpu()).all() and not back[n:].any()
    want_root = tree.root()
    sub_layer = tree.commitment.layers[3].toCpu().tobytes()          # the 8 subtree roots of the single tree
    del tree, back
    # (d) sharded path, all "ranks" on this device: rank r transforms columns [32 r, 32 r + 32) and stores into all 8 buffers
    rows, lc = (1 << ext) // G, n_cols // G
    shards = [tb.CudaColumn.uninitialized(n_cols * rows) for _ in range(G)]
    work = tb.CudaColumn.uninitialized(lc << ext)
    bases = _lib.dptr_array([s.ptr for s in shards])
    for r in range(G):
        _lib.check(L.tstwo_lde_columns_to_peers_overlapped(dtrace.ptr + 4 * n * lc * r, lc, log, blow, tw.handle, 0, work.ptr, bases, G,
                                                           r * lc, 0, 4 if r % 2 else 1, 62 if r % 2 else 0))
    roots = tb.CudaColumn.uninitialized(8 * G)
    sub = tb.CudaColumn.uninitialized(8 * ((2 * rows) - 1))
    logs = (C.c_int * n_cols)(*([ext - 3] * n_cols))
    for d in range(G):
        ptrs = _lib.dptr_array([shards[d].ptr + 4 * rows * c for c in range(n_cols)])
        _lib.check(L.tstwo_merkle_commit(ptrs, logs, n_cols, sub.ptr, None))
        _lib.check(L.tstwo_copy(roots.ptr + 32 * d, sub.ptr, 32))
    assert roots.toCpu().tobytes() == sub_layer, "subtree roots differ from layer 3 of the single tree"
    prev = roots
    for k in (2, 1, 0):
        cur = tb.CudaColumn.uninitialized(8 << k)
        _lib.check(L.tstwo_merkle_commit_layer(k, prev.ptr, None, 0, cur.ptr))
        prev = cur
    assert prev.toCpu().tobytes() == want_root


def test_shutdown_releases_side_streams_and_reinit_works():
    """tstwo_shutdown frees what the pipeline keeps between calls (coefficient scratch, the side stream and events of the
    overlapped exchange, the scatter streams); a second tstwo_init in the same process must give the same results.  Run in a
    child process: shutting the library down under the other tests' columns would pull the memory from under them."""
    import os, subprocess, sys, textwrap
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = textwrap.dedent("""
        import ctypes as C, numpy as np, sys
        sys.path.insert(0, %r)
        from tstwo_b200 import _lib
        P = 2147483647
        log, blow, n_cols, G = 15, 1, 6, 2
        ext, rows = log + blow, (1 << (log + blow)) // G
        trace = np.random.default_rng(5).integers(0, P, n_cols << log, dtype=np.uint32)
        outs = []
        for rep in range(2):
            L = _lib.init(0)
            chk = _lib.check
            def dmalloc(words):
                p = C.c_uint64(); chk(L.tstwo_malloc(4 * words, C.byref(p))); return p.value
            tw = C.c_uint64(); chk(L.tstwo_twiddles_precompute(1 << (31 - (ext - 1) - 2), ext - 1, C.byref(tw)))
            d_trace, work = dmalloc(n_cols << log), dmalloc(n_cols << ext)
            shards = [dmalloc(n_cols * rows) for _ in range(G)]
            chk(L.tstwo_upload(d_trace, trace.ctypes.data, 4 * trace.size))
            chk(L.tstwo_lde_columns_to_peers_overlapped(d_trace, n_cols, log, blow, tw.value, 0, work, _lib.dptr_array(shards), G, 0, 0, 3, 50))
            sharded = dmalloc(n_cols << ext)
            chk(L.tstwo_lde_columns_sharded(d_trace, n_cols, log, blow, tw.value, 0, sharded, G, 0))
            copies = [dmalloc(n_cols * rows) for _ in range(G)]
            chk(L.tstwo_scatter_shards(sharded, n_cols, rows, _lib.dptr_array(copies), G, 0)); chk(L.tstwo_scatter_join())
            got = []
            for a, b in zip(shards, copies):
                x, y = np.empty(n_cols * rows, dtype=np.uint32), np.empty(n_cols * rows, dtype=np.uint32)
                chk(L.tstwo_download(x.ctypes.data, a, 4 * x.size)); chk(L.tstwo_download(y.ctypes.data, b, 4 * y.size))
                assert (x == y).all(), "peer stores and copy-engine scatter disagree"
                got.append(x)
            outs.append(np.concatenate(got))
            assert L.tstwo_shutdown() == 0
            _lib._inited = None
        assert (outs[0] == outs[1]).all()
        print("reinit ok")
    """ % root)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "reinit ok" in r.stdout, r.stdout + r.stderr


def test_peer_flags_signal_wait_put(tb):
    """tstwo_peer_signal / _wait / _put on one device (the flag block and the "peer" are local memory): the wait returns once
    every flag has reached the value, in wrap-around order, and put delivers its words before its flag; twiddle trees freed and
    asked for again come back without being rebuilt (spare list) with identical tables."""
    from tstwo_b200 import _lib
    L, chk = _lib.load(), _lib.check
    blk = tb.CudaColumn.zeros(64)
    for value in (1, 2, 0xFFFFFFFF, 3):                       # 0xFFFFFFFF -> 3 wraps: (int32)(3 - 0xFFFFFFFF) = 4 >= 0
        chk(L.tstwo_peer_signal(_lib.dptr_array([blk.ptr + 4 * i for i in range(5)]), 5, value))
        chk(L.tstwo_peer_wait(blk.ptr, 5, value, 2000))
        chk(L.tstwo_sync())
        assert list(blk.toCpu()[:6]) == [value] * 5 + [0]
    chk(L.tstwo_peer_wait(blk.ptr, 5, 2, 2000))               # already past 2 (wrap-around order): returns at once
    src = tb.CudaColumn.from_cpu(np.arange(100, 108, dtype=np.uint32))
    chk(L.tstwo_peer_put(blk.ptr + 4 * 16, src.ptr, 8, blk.ptr + 4 * 40, 9))
    chk(L.tstwo_peer_wait(blk.ptr + 4 * 40, 1, 9, 2000))
    chk(L.tstwo_sync())
    got = blk.toCpu()
    assert list(got[16:24]) == list(range(100, 108)) and got[40] == 9
    assert L.tstwo_peer_signal(None, 1, 1) != 0 and L.tstwo_peer_wait(blk.ptr, 65, 1, 0) != 0
    # spare twiddle trees
    coset = tb.CanonicCoset.new(12).half_coset()
    t1 = tb.CudaBackend.precomputeTwiddles(coset)
    a = t1.twiddles.toCpu().copy()
    n0 = L.tstwo_launch_count()
    del t1
    t2 = tb.CudaBackend.precomputeTwiddles(coset)
    assert L.tstwo_launch_count() == n0, "a freed tree of the same coset is reused, not rebuilt"
    assert (t2.twiddles.toCpu() == a).all()


def test_lde_columns_to_peers_log23_uses_the_nine_layer_pass(tb):
    """The multi-GPU store path at log 23 (contiguous pass + the 512-row pass, the latter on the main stream, the storing pass on the
    side stream): row shards equal the single-device evaluations of tstwo_lde_commit."""
    import ctypes as C
    from tstwo_b200 import _lib
    L = _lib.init(0)
    log, blow, n_cols, n_shards = 23, 1, 3, 4
    ext = log + blow
    rows = (1 << ext) // n_shards
    rng = np.random.default_rng(23)
    trace = rng.integers(0, P, (n_cols, 1 << log), dtype=np.uint32)
    tw = tb.CudaBackend.precomputeTwiddles(tb.CanonicCoset.new(ext).half_coset())
    dtrace = tb.CudaColumn.from_cpu(trace.reshape(-1))
    evals = tb.CudaColumn.uninitialized(n_cols << ext)
    root = (C.c_uint8 * 32)()
    _lib.check(L.tstwo_lde_commit(dtrace.ptr, n_cols, log, blow, tw.handle, 0, evals.ptr, 0, root, 0))
    want = evals.toCpu().reshape(n_cols, 1 << ext)
    for groups, peer_sms in ((1, 0), (3, 62)):
        shards = [tb.CudaColumn.zeros(n_cols * rows) for _ in range(n_shards)]
        work = tb.CudaColumn.uninitialized(n_cols << ext)
        _lib.check(L.tstwo_lde_columns_to_peers_overlapped(dtrace.ptr, n_cols, log, blow, tw.handle, 0, work.ptr,
                                                           _lib.dptr_array([s.ptr for s in shards]), n_shards, 0, 0, groups, peer_sms))
        for d in range(n_shards):
            got = shards[d].toCpu().reshape(n_cols, rows)
            assert (got == want[:, d * rows:(d + 1) * rows]).all(), f"groups {groups} shard {d}"


@pytest.mark.parametrize("log,n_cols,cuts", [(12, 64, [32]), (10, 48, [16, 32]), (13, 40, [16]), (9, 33, [32]), (7, 64, [32]), (12, 16, []),
                                             (16, 256, [128]), (11, 100, [64, 96])])
def test_merkle_leaf_absorb_in_column_chunks_equals_the_one_shot_tree(tb, log, n_cols, cuts):
    """tstwo_merkle_leaf_absorb over column chunks + tstwo_merkle_commit_nodes = tstwo_merkle_commit over all columns (every layer),
    and the leaf layer equals hashlib's BLAKE2s of the rows; chunks of >= 256 rows take the bulk-copy kernel, smaller ones and
    columns that are not 16-byte spaced the per-thread one (log 7; the odd stride below)."""
    from tstwo_b200 import _lib
    L, chk = _lib.init(0), _lib.check
    n = 1 << log
    rng = np.random.default_rng(1000 * log + n_cols)
    for stride in (n, n + 4, n + 1):
        data = rng.integers(0, P, (n_cols, stride), dtype=np.uint32)
        block = tb.CudaColumn.from_cpu(data.reshape(-1))
        want_layers = tb.CudaColumn.uninitialized(8 * ((2 << log) - 1))
        ptrs = _lib.dptr_array([block.ptr + 4 * stride * c for c in range(n_cols)])
        chk(L.tstwo_merkle_commit(ptrs, (C.c_int * n_cols)(*([log] * n_cols)), n_cols, want_layers.ptr, None))
        layers = tb.CudaColumn.from_cpu(np.full(8 * ((2 << log) - 1), 0x2B2B2B2B, dtype=np.uint32))
        edges = [0] + cuts + [n_cols]
        for c0, c1 in zip(edges[:-1], edges[1:]):
            chk(L.tstwo_merkle_leaf_absorb(block.ptr + 4 * stride * c0, stride, c0, c1 - c0, n_cols, log, layers.ptr))
        root = (C.c_uint8 * 32)()
        chk(L.tstwo_merkle_commit_nodes(log, layers.ptr, root))
        got, want = layers.toCpu(), want_layers.toCpu()
        assert (got == want).all(), f"stride {stride}"
        assert bytes(root) == want[:8].tobytes()
        for i in (0, 1, n // 2 + 3 if n > 8 else 0, n - 1):
            leaf = got[8 * ((1 << log) - 1 + i): 8 * ((1 << log) + i)].tobytes()
            assert leaf == hashlib.blake2s(data[:, i].astype("<u4").tobytes()).digest()
    blk = tb.CudaColumn.uninitialized(n_cols * n)
    lay = tb.CudaColumn.uninitialized(8 * ((2 << log) - 1))
    assert L.tstwo_merkle_leaf_absorb(blk.ptr, n, 8, 8, n_cols, log, lay.ptr) != 0 and b"block aligned" in L.tstwo_last_error()
    assert L.tstwo_merkle_leaf_absorb(blk.ptr, n, 0, n_cols + 16, n_cols, log, lay.ptr) != 0


@pytest.mark.parametrize("groups,peer_sms,n_cols,n_shards", [(2, 62, 4, 2), (4, 66, 8, 4), (4, 40, 16, 1), (8, 62, 16, 2)])
def test_lde_columns_to_peers_halves(tb, orc, groups, peer_sms, n_cols, n_shards):
    """tstwo_lde_columns_to_peers_halves on local shard buffers: the first half of the rank's columns lands at col_offset_lo.., the
    second at col_offset_hi.., nothing else is touched, both flag sets carry the value once the side stream is joined, and the call
    leaves the side stream unjoined without losing work (repeated)."""
    from tstwo_b200 import _lib
    L, chk = _lib.init(0), _lib.check
    log, blow = 15, 1
    ext = log + blow
    rows = (1 << ext) // n_shards
    lo, hi, total = 1, n_cols + 3, 2 * n_cols + 4
    rng = np.random.default_rng(100 * groups + n_cols)
    trace = rng.integers(0, P, (n_cols, 1 << log), dtype=np.uint32)
    tw = tb.CudaBackend.precomputeTwiddles(tb.CanonicCoset.new(ext).half_coset())
    root_coset = tb.CanonicCoset.new(ext).half_coset()
    otw, oitw = orc.precompute_twiddles(root_coset.initial_index.value, root_coset.log_size)
    _, evals, _ = orc.lde_commit(trace, blow, otw, oitw, want_coeffs=True)
    dtrace = tb.CudaColumn.from_cpu(trace.reshape(-1))
    flags = tb.CudaColumn.zeros(64)
    half = _lib.dptr_array([flags.ptr + 4 * i for i in range(3)])
    arrive = _lib.dptr_array([flags.ptr + 4 * (8 + i) for i in range(3)])
    where = {lo + c: c for c in range(n_cols // 2)}
    where.update({hi + c: n_cols // 2 + c for c in range(n_cols // 2)})
    for rep in range(1, 4):
        shards = [tb.CudaColumn.from_cpu(np.full(total * rows, 7, dtype=np.uint32)) for _ in range(n_shards)]
        work = tb.CudaColumn.uninitialized(n_cols << ext)
        chk(L.tstwo_lde_columns_to_peers_halves(dtrace.ptr, n_cols, log, blow, tw.handle, 0, work.ptr,
                                                _lib.dptr_array([s.ptr for s in shards]), n_shards, lo, hi, 0, groups, peer_sms,
                                                half, arrive, 3, rep))
        chk(L.tstwo_peer_wait(flags.ptr, 3, rep, 5000))          # half flags: on the library's stream, beside the side stream
        chk(L.tstwo_peer_join())
        chk(L.tstwo_sync())
        f = flags.toCpu()
        assert list(f[:3]) == [rep] * 3 and list(f[8:11]) == [rep] * 3 and f[3] == 0 and f[11] == 0
        for d in range(n_shards):
            got = shards[d].toCpu().reshape(total, rows)
            for c in range(total):
                if c in where:
                    assert (got[c] == evals[where[c]][d * rows:(d + 1) * rows]).all(), f"rep {rep} shard {d} column {c}"
                else:
                    assert (got[c] == 7).all(), f"shard {d} column {c} must be untouched"
    buf = tb.CudaColumn.uninitialized(4 << ext)
    bases = _lib.dptr_array([buf.ptr])
    for bad in ((3, 62), (1, 62), (2, 0)):                       # odd group count, one group, no side SMs
        rc = L.tstwo_lde_columns_to_peers_halves(dtrace.ptr, 4, log, blow, tw.handle, 0, work.ptr, bases, 1, 0, 2, 0, bad[0], bad[1], half, arrive, 3, 9)
        assert rc != 0 and b"halves" in L.tstwo_last_error(), bad
    chk(L.tstwo_peer_join())

"""Parity of the CUDA Blake2s Merkle commit with the oracle / hashlib.  Mirrors
packages/core/test/backend/cpu/blake2.test.ts:55-176 and test/vcs/blake2_merkle.test.ts inputs (LCG seed 0)."""
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


@pytest.mark.parametrize("n_cols", [0, 1, 3, 15, 16, 17, 31, 32, 33, 48, 100, 256])
def test_commit_layer(tb, orc, n_cols):
    rng = np.random.default_rng(n_cols)
    ops = tb.CudaBackend.merkleOps
    for log in [0, 1, 5, 9]:
        cols = [rng.integers(0, P, 1 << log, dtype=np.uint32) for _ in range(n_cols)]
        dcols = [tb.CudaColumn.from_cpu(c) for c in cols]
        leaf = ops.commitOnLayer(log, None, dcols)
        exp = orc.merkle_commit_layer(log, None, cols)
        got = leaf.toCpu().view(np.uint8).reshape(-1, 32)
        assert (got == exp).all()
        i = (1 << log) - 1
        assert got[i].tobytes() == hashlib.blake2s(b"".join(struct.pack("<I", int(c[i])) for c in cols)).digest()
        if log > 0:
            k = n_cols % 5
            up = ops.commitOnLayer(log - 1, leaf, [tb.CudaColumn.from_cpu(c[: 1 << (log - 1)]) for c in cols[:k]])
            exp_up = orc.merkle_commit_layer(log - 1, exp, [c[: 1 << (log - 1)] for c in cols[:k]])
            assert (up.toCpu().view(np.uint8).reshape(-1, 32) == exp_up).all()


def test_commit_contiguous_block(tb, orc):
    rng = np.random.default_rng(9)
    n_cols, log = 40, 10
    block = rng.integers(0, P, (n_cols, 1 << log), dtype=np.uint32)
    d = tb.CudaColumn.from_cpu(block.reshape(-1))
    mp = tb.MerkleProver.commit(tb.CudaBackend.merkleOps, [d.slice(c << log, 1 << log) for c in range(n_cols)])
    exp = orc.merkle_commit(list(block))
    assert mp.root() == exp[0][0].tobytes()
    for k in range(log + 1):
        assert (mp.layer_bytes(k) == exp[k]).all()


def test_commit_mixed_sizes_lcg(tb, orc):
    seed = 0
    def nxt():
        nonlocal seed
        seed = (1664525 * seed + 1013904223) % 2**32
        return seed
    logs = [3 + nxt() % 2 for _ in range(10)]
    cols = [np.array([nxt() % (1 << 30) for _ in range(1 << l)], dtype=np.uint32) for l in logs]
    mp = tb.MerkleProver.commit(tb.CudaBackend.merkleOps, [tb.CudaColumn.from_cpu(c) for c in cols])
    exp = orc.merkle_commit(cols)
    assert len(mp.layers) == len(exp)
    for k in range(len(exp)):
        assert (mp.layer_bytes(k) == exp[k]).all()


def test_commit_empty(tb):
    mp = tb.MerkleProver.commit(tb.CudaBackend.merkleOps, [])
    assert mp.root().hex() == "69217a3079908094e11121d042354a7c1f55b6482ca1a51e1b250dfd1ed0eef9"


@pytest.mark.parametrize("n_cols,log", [(1, 8), (7, 8), (16, 9), (17, 10), (33, 8), (48, 11), (100, 9), (255, 8), (256, 10), (300, 8)])
def test_leaf_layer_bulk_copy_kernel(tb, orc, n_cols, log):
    """Leaf layers of uniformly spaced columns with >= 256 rows go through merkle_leaf_tma_kernel (message blocks arrive as
    [16 columns x 256 rows] tensor boxes, the ragged last block zero-filled by the copy engine): same digests as the per-thread
    load kernel and as the oracle (vcs/blake2_merkle.ts:9-24)."""
    from tstwo_b200._lib import load
    L = load()
    rng = np.random.default_rng(1000 * n_cols + log)
    block = rng.integers(0, P, (n_cols, 1 << log), dtype=np.uint32)
    d = tb.CudaColumn.from_cpu(block.reshape(-1))
    cols = [d.slice(c << log, 1 << log) for c in range(n_cols)]
    ops = tb.CudaBackend.merkleOps
    exp = orc.merkle_commit_layer(log, None, list(block))
    try:
        got = {}
        for mode in (1, 0):
            L.tstwo_debug_set_leaf_mode(mode)
            got[mode] = ops.commitOnLayer(log, None, cols).toCpu().view(np.uint8).reshape(-1, 32)
            assert L.tstwo_debug_last_leaf_mode() == mode
    finally:
        L.tstwo_debug_set_leaf_mode(1)
    assert (got[1] == exp).all()
    assert (got[0] == exp).all()


def test_leaf_layer_scattered_columns_fall_back(tb, orc):
    """Columns that are separate allocations (an address table) keep the per-thread load kernel."""
    from tstwo_b200._lib import load
    L = load()
    rng = np.random.default_rng(77)
    cols = [rng.integers(0, P, 1 << 9, dtype=np.uint32) for _ in range(20)]
    pad = [tb.CudaColumn.uninitialized(4 * (c + 1)) for c in range(20)]   # break any accidental uniform spacing
    dcols = [tb.CudaColumn.from_cpu(c) for c in cols]
    got = tb.CudaBackend.merkleOps.commitOnLayer(9, None, dcols).toCpu().view(np.uint8).reshape(-1, 32)
    assert (got == orc.merkle_commit_layer(9, None, cols)).all()
    del pad

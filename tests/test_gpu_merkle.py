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

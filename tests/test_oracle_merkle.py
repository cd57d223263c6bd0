"""Blake2s + Merkle commit of the oracle vs the reference's literals and Python hashlib (independent
RFC 7693 implementation).  Ports packages/core/test/vcs/blake2_hash.test.ts:7-8,127-128,166,
test/backend/cpu/blake2.test.ts:55-176 and the prepareMerkle input generator (src/vcs/test_utils.ts:46-144)."""
import hashlib
import struct

import numpy as np


def h(b):
    return hashlib.blake2s(b).digest()


def test_blake2s_kats(orc):
    assert orc.blake2s(b"").hex() == "69217a3079908094e11121d042354a7c1f55b6482ca1a51e1b250dfd1ed0eef9"
    assert orc.blake2s(b"a").hex() == "4a0d129873403037c2cd9b9048203687f6233fb6738956e0349bd4320fec3e90"
    assert orc.blake2s(orc.blake2s(b"a") + orc.blake2s(b"b")).hex() == \
        "2d12d4f7a2c2c9e02fc6300b0d23c772457aa5e30d1d69e7b589b8a48afe5425"
    rng = np.random.default_rng(0)
    for n in [1, 3, 4, 63, 64, 65, 127, 128, 129, 1024, 1025, 4000]:
        m = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
        assert orc.blake2s(m) == h(m)


def py_commit(cols):
    """MerkleProver.commit with hashNode semantics (vcs/prover.ts:13-30, vcs/blake2_merkle.ts:9-24)."""
    if not cols:
        return [[h(b"")]]
    cols = sorted(cols, key=lambda c: -len(c))
    max_log = len(cols[0]).bit_length() - 1
    layers = []
    for log in range(max_log, -1, -1):
        lc = [c for c in cols if len(c) == 1 << log]
        prev = layers[-1] if layers else None
        layer = []
        for i in range(1 << log):
            msg = b""
            if prev is not None:
                msg += prev[2 * i] + prev[2 * i + 1]
            msg += b"".join(struct.pack("<I", int(c[i])) for c in lc)
            layer.append(h(msg))
        layers.append(layer)
    layers.reverse()
    return layers


def lcg_cols():
    seed = 0
    def nxt():
        nonlocal seed
        seed = (1664525 * seed + 1013904223) % 2**32
        return seed
    logs = [3 + nxt() % 2 for _ in range(10)]
    return [[nxt() % (1 << 30) for _ in range(1 << l)] for l in logs]


def test_commit_layer_leaf_and_node(orc):
    rng = np.random.default_rng(1)
    for n_cols in [0, 1, 3, 15, 16, 17, 31, 32, 33, 100]:
        for log in [0, 1, 4]:
            cols = [rng.integers(0, orc.P, 1 << log, dtype=np.uint32) for _ in range(n_cols)]
            leaf = orc.merkle_commit_layer(log, None, cols)
            for i in range(1 << log):
                assert leaf[i].tobytes() == h(b"".join(struct.pack("<I", int(c[i])) for c in cols))
            if log > 0:
                up_cols = [rng.integers(0, orc.P, 1 << (log - 1), dtype=np.uint32) for _ in range(n_cols % 3)]
                up = orc.merkle_commit_layer(log - 1, leaf, up_cols)
                for i in range(1 << (log - 1)):
                    msg = leaf[2 * i].tobytes() + leaf[2 * i + 1].tobytes() + b"".join(struct.pack("<I", int(c[i])) for c in up_cols)
                    assert up[i].tobytes() == h(msg)


def test_commit_mixed_sizes_lcg(orc):
    cols = lcg_cols()
    got = orc.merkle_commit(cols)
    exp = py_commit(cols)
    assert len(got) == len(exp)
    for g, e in zip(got, exp):
        assert [x.tobytes() for x in g] == e


def test_commit_uniform_and_empty(orc):
    rng = np.random.default_rng(2)
    cols = [rng.integers(0, orc.P, 64, dtype=np.uint32) for _ in range(20)]
    got = orc.merkle_commit(cols)
    exp = py_commit([c.tolist() for c in cols])
    assert got[0][0].tobytes() == exp[0][0]
    assert [x.tobytes() for x in got[6]] == exp[6]
    assert orc.merkle_commit([])[0][0].tobytes() == h(b"")

"""MerkleProver.decommit / MerkleVerifier.verify restated in the oracle: the reference's own scenario
(core/src/vcs/test_utils.ts:78-144, LCG seed 0: 10 columns of log size 3 or 4, 3 queries per size) and its negative
tests (core/test/vcs/blake2_merkle.test.ts:11-99)."""
import numpy as np
import pytest


class Lcg:
    def __init__(self, seed=0): self.s = seed
    def next(self):
        self.s = (1664525 * self.s + 1013904223) % 2**32
        return self.s
    def rng(self, lo, hi): return lo + self.next() % (hi - lo)


def prepare(orc):
    r = Lcg(0)
    logs = [r.rng(3, 5) for _ in range(10)]
    cols = [np.array([r.rng(0, 1 << 30) for _ in range(1 << l)], dtype=np.uint32) for l in logs]
    layers = orc.merkle_commit(cols)
    queries = {}
    for l in (4, 3):
        queries[l] = sorted({r.rng(0, 1 << l) for _ in range(3)})
    return logs, cols, layers, queries


def test_decommit_verifies(orc):
    logs, cols, layers, queries = prepare(orc)
    queried, hw, cw = orc.merkle_decommit(layers, queries, cols)
    orc.merkle_verify(layers[0][0].tobytes(), logs, queries, queried, hw, cw)
    # every queried value is the column value at the queried position, larger columns first
    exp = []
    for l in (4, 3):
        for q in queries[l]:
            exp += [int(c[q]) for c, cl in zip(cols, logs) if cl == l]
    assert sorted(queried) == sorted(exp) and len(queried) == len(exp)


def test_negative_cases(orc):
    logs, cols, layers, queries = prepare(orc)
    root = layers[0][0].tobytes()
    queried, hw, cw = orc.merkle_decommit(layers, queries, cols)
    bad = list(queried); bad[3] = (bad[3] + 1) % (2**31 - 1)
    with pytest.raises(ValueError, match="RootMismatch"):
        orc.merkle_verify(root, logs, queries, bad, hw, cw)
    badh = list(hw); badh[4] = bytes(32)
    with pytest.raises(ValueError, match="RootMismatch"):
        orc.merkle_verify(root, logs, queries, queried, badh, cw)
    with pytest.raises(ValueError, match="WitnessTooShort"):
        orc.merkle_verify(root, logs, queries, queried, hw[:-1], cw)
    with pytest.raises(ValueError, match="WitnessTooLong"):
        orc.merkle_verify(root, logs, queries, queried, hw + [bytes(32)], cw)
    with pytest.raises(ValueError, match="WitnessTooLong"):
        orc.merkle_verify(root, logs, queries, queried, hw, cw + [0])
    with pytest.raises(ValueError, match="TooFewQueriedValues"):
        orc.merkle_verify(root, logs, queries, queried[:-1], hw, cw)
    with pytest.raises(ValueError, match="TooManyQueriedValues"):
        orc.merkle_verify(root, logs, queries, queried + [0], hw, cw)


def test_uniform_tree_path_matches_hashlib(orc):
    import hashlib
    rng = np.random.default_rng(4)
    cols = [rng.integers(0, 2**31 - 1, 64, dtype=np.uint32) for _ in range(5)]
    layers = orc.merkle_commit(cols)
    queried, hw, cw = orc.merkle_decommit(layers, {6: [37]}, cols)
    assert queried == [int(c[37]) for c in cols] and cw == [] and len(hw) == 6
    h = hashlib.blake2s(b"".join(int(v).to_bytes(4, "little") for v in queried)).digest()
    idx = 37
    for k in range(6):
        sib = hw[k]
        h = hashlib.blake2s(h + sib if idx % 2 == 0 else sib + h).digest()
        idx //= 2
    assert h == layers[0][0].tobytes()

"""The callers of the hot path: MerkleProver.decommit on device trees, TreeBuilder /
CommitmentSchemeProver with the Fiat-Shamir channel.  Mirrors core/test/vcs/blake2_merkle.test.ts (LCG scenario of
vcs/test_utils.ts:78-144) and the commit flow of the Rust text in core/src/pcs/prover.ts:28-84,170-235."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
P = 2147483647


@pytest.fixture(scope="module")
def tb():
    import tstwo_b200
    return tstwo_b200


class Lcg:
    def __init__(self, seed=0): self.s = seed
    def next(self):
        self.s = (1664525 * self.s + 1013904223) % 2**32
        return self.s
    def rng(self, lo, hi): return lo + self.next() % (hi - lo)


def test_decommit_lcg_scenario(tb, orc):
    r = Lcg(0)
    logs = [r.rng(3, 5) for _ in range(10)]
    cols = [np.array([r.rng(0, 1 << 30) for _ in range(1 << l)], dtype=np.uint32) for l in logs]
    queries = {l: sorted({r.rng(0, 1 << l) for _ in range(3)}) for l in (4, 3)}
    dcols = [tb.CudaColumn.from_cpu(c) for c in cols]
    mp = tb.MerkleProver.commit(tb.CudaBackend.merkleOps, dcols)
    values, dec = mp.decommit(queries, dcols)
    layers = orc.merkle_commit(cols)
    q, hw, cw = orc.merkle_decommit(layers, queries, cols)
    assert list(values) == q and dec.hashWitness == hw and list(dec.columnWitness) == cw
    orc.merkle_verify(mp.root(), logs, queries, list(values), dec.hashWitness, list(dec.columnWitness))


@pytest.mark.parametrize("n_cols,log,n_queries", [(1, 10, 1), (7, 12, 5), (40, 14, 30), (256, 10, 64)])
def test_decommit_uniform_trees(tb, orc, n_cols, log, n_queries):
    rng = np.random.default_rng(n_cols * 100 + log)
    block = rng.integers(0, P, (n_cols, 1 << log), dtype=np.uint32)
    d = tb.CudaColumn.from_cpu(block.reshape(-1))
    dcols = [d.slice(c << log, 1 << log) for c in range(n_cols)]
    mp = tb.MerkleProver.commit(tb.CudaBackend.merkleOps, dcols)
    queries = {log: sorted({int(x) for x in rng.integers(0, 1 << log, n_queries)})}
    values, dec = mp.decommit(queries, dcols)
    assert (values.reshape(-1, n_cols) == block[:, queries[log]].T).all() and len(dec.columnWitness) == 0
    orc.merkle_verify(mp.root(), [log] * n_cols, queries, list(values), dec.hashWitness, [])
    with pytest.raises(ValueError, match="RootMismatch"):
        orc.merkle_verify(mp.root(), [log] * n_cols, queries, [(int(values[0]) + 1) % P] + list(values[1:]), dec.hashWitness, [])
    with pytest.raises(IndexError):
        mp.decommit({log: [1 << log]}, dcols)


def test_commitment_scheme_with_channel(tb, orc):
    """Two trees (the second with two polynomial sizes) through TreeBuilder.extend_evals / extend_polys / commit: roots,
    transcript and decommitments against the oracle."""
    from tstwo_b200.channel import Blake2sChannel
    rng = np.random.default_rng(11)
    blow = 1
    tw = tb.CudaBackend.precomputeTwiddles(tb.CanonicCoset.new(11 + blow).halfCoset())
    root_init, root_log = orc.canonic_half_coset(11 + blow)
    otw, oitw = orc.precompute_twiddles(root_init, root_log)
    channel, ochannel = Blake2sChannel.create(), orc.Channel()
    scheme = tb.CommitmentSchemeProver(tb.PcsConfig(5, tb.FriConfig(0, blow, 3)), tw)

    def host_tree(cols_by_log):
        """oracle: interpolate (trace given as evaluations) or take coefficients, evaluate on the blown-up domain, commit"""
        ev_cols, logs = [], []
        for log, arrs, are_evals in cols_by_log:
            h_init, _ = orc.canonic_half_coset(log)
            e_init, _ = orc.canonic_half_coset(log + blow)
            for a in arrs:
                coeffs = orc.interpolate(a, h_init, (root_init, root_log), oitw) if are_evals else a
                ev_cols.append(orc.evaluate(coeffs, log + blow, e_init, (root_init, root_log), otw))
                logs.append(log + blow)
        return ev_cols, logs, orc.merkle_commit(ev_cols)

    # tree 0: 6 trace columns of log 10, given as evaluations
    t0 = [rng.integers(0, P, 1 << 10, dtype=np.uint32) for _ in range(6)]
    dom10 = tb.CanonicCoset.new(10).circleDomain()
    b0 = scheme.tree_builder()
    span = b0.extend_evals([tb.CudaCircleEvaluation.new(dom10, c) for c in t0])
    assert span == tb.TreeSubspan(0, 0, 6)
    b0.commit(channel)
    ev0, logs0, layers0 = host_tree([(10, t0, True)])
    ochannel.mix_root(layers0[0][0].tobytes())
    assert scheme.roots() == [layers0[0][0].tobytes()] and channel.digest() == ochannel.digest

    # tree 1: 3 polynomials of log 8 and 2 of log 11, given as coefficients, interleaved
    p8 = [rng.integers(0, P, 1 << 8, dtype=np.uint32) for _ in range(3)]
    p11 = [rng.integers(0, P, 1 << 11, dtype=np.uint32) for _ in range(2)]
    b1 = scheme.tree_builder()
    assert b1.extend_polys([tb.CudaCirclePoly.new(p8[0]), tb.CudaCirclePoly.new(p11[0])]) == tb.TreeSubspan(1, 0, 2)
    assert b1.extend_polys([tb.CudaCirclePoly.new(p8[1]), tb.CudaCirclePoly.new(p8[2]), tb.CudaCirclePoly.new(p11[1])]) == tb.TreeSubspan(1, 2, 5)
    b1.commit(channel)
    order = [(8, p8[0]), (11, p11[0]), (8, p8[1]), (8, p8[2]), (11, p11[1])]
    ev1, logs1 = [], []
    for log, a in order:
        e, l, _ = host_tree([(log, [a], False)])
        ev1 += e; logs1 += l
    layers1 = orc.merkle_commit(ev1)
    ochannel.mix_root(layers1[0][0].tobytes())
    assert scheme.roots()[1] == layers1[0][0].tobytes() and channel.digest() == ochannel.digest
    assert len(scheme.polynomials()[1]) == 5 and len(scheme.evaluations()[0]) == 6
    assert channel.draw_felt() == ochannel.draw_felt()

    # decommit both trees at query positions per LDE log size, verify against the roots
    queries = {12: [5, 77, 4000], 11: [0, 1023, 2047], 9: [3, 300]}
    res = scheme.decommit(queries)
    for (values, dec), logs, evs, layers in zip(res, (logs0, logs1), (ev0, ev1), (layers0, layers1)):
        q, hw, cw = orc.merkle_decommit(layers, queries, evs)
        assert list(values) == q and dec.hashWitness == hw and list(dec.columnWitness) == cw
        orc.merkle_verify(layers[0][0].tobytes(), logs, queries, list(values), dec.hashWitness, list(dec.columnWitness))

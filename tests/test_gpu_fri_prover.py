"""FriProver.commit / decommit on the device (tstwo_b200/fri.py) against the oracle's restatement of the same flow:
layer roots, folding alphas (through the shared transcript), last-layer polynomial, FRI witnesses and Merkle
decommitments; every layer's decommitment is then checked with the oracle's MerkleVerifier.  The polynomial inputs are
low degree so that the last-layer degree check (core/src/fri.ts:733-748) is exercised for real."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
P = 2147483647


@pytest.fixture(scope="module")
def tb():
    import tstwo_b200
    return tstwo_b200


def secure_column_of_degree(tb, orc, log_degree, log_blowup, seed):
    """4 coordinate polynomials with 2^log_degree coefficients, evaluated on CanonicCoset(log_degree + log_blowup)."""
    rng = np.random.default_rng(seed)
    log = log_degree + log_blowup
    root_init, root_log = orc.canonic_half_coset(log)
    tw, _ = orc.precompute_twiddles(root_init, root_log)
    coords = []
    for _ in range(4):
        coeffs = rng.integers(0, P, 1 << log_degree, dtype=np.uint32)
        coords.append(orc.evaluate(coeffs, log, root_init, (root_init, root_log), tw))
    return log, coords


@pytest.mark.parametrize("log_degrees,log_last,blow,n_queries", [((6,), 0, 1, 3), ((8, 6, 5), 2, 1, 5), ((7, 4), 0, 2, 4), ((10,), 3, 1, 8)])
def test_fri_commit_and_decommit(tb, orc, log_degrees, log_last, blow, n_queries):
    from tstwo_b200.channel import Blake2sChannel
    from tstwo_b200.fri import FriProver, Queries
    cols = [secure_column_of_degree(tb, orc, d, blow, 100 * d + blow) for d in log_degrees]
    channel, ochannel = Blake2sChannel.create(), orc.Channel()
    channel.mix_u64(42); ochannel.mix_u64(42)
    dev_cols = [tb.SecureEvaluation(tb.CanonicCoset.new(log).circleDomain(), tb.SecureColumnByCoords.from_cpu(c)) for log, c in cols]
    config = tb.FriConfig(log_last, blow, n_queries)
    prover = FriProver.commit(channel, config, dev_cols)
    state = orc.fri_commit(ochannel, log_last, blow, cols)
    assert prover.first_layer.merkle_tree.root() == state["first_layers"][0][0].tobytes()
    assert len(prover.inner_layers) == len(state["inner"]) == log_degrees[0] + blow - 1 - (log_last + blow)
    for layer, ol in zip(prover.inner_layers, state["inner"]):
        assert layer.merkle_tree.root() == ol["layers"][0][0].tobytes()
        assert all((a == b).all() for a, b in zip(layer.evaluation.values.to_cpu(), ol["evaluation"]))
    assert prover.last_layer_poly.coeffs == state["last_poly"] and len(prover.last_layer_poly) == 1 << log_last
    assert channel.digest() == ochannel.digest

    # decommit: queries from the transcript, proof layer by layer
    proof, by_log = prover.decommit(channel)
    max_log = cols[0][0]
    oq = Queries.generate(ochannel, max_log, n_queries)
    assert by_log == {log: oq.fold(max_log - log).positions for log, _ in cols}
    exp = orc.fri_decommit_on_queries(state, oq.positions, max_log)
    got = [proof.first_layer] + proof.inner_layers
    assert len(got) == len(exp)
    for g, (w, (queried, hw, cw), root) in zip(got, exp):
        assert g.fri_witness == w and g.commitment == root
        assert g.decommitment.hashWitness == hw and list(g.decommitment.columnWitness) == cw
    # every layer's decommitment verifies against its root
    pos_by_log = {}
    for log, _ in cols:
        pos_by_log[log] = orc.fri_decommitment_positions(oq.fold(max_log - log).positions, 1)[0]
    first_logs = [log for log, _ in cols for _ in range(4)]
    queried, hw, cw = exp[0][1]
    orc.merkle_verify(exp[0][2], first_logs, pos_by_log, queried, hw, cw)
    q = oq.fold(1).positions
    for ol, (w, (queried, hw, cw), root) in zip(state["inner"], exp[1:]):
        pos = orc.fri_decommitment_positions(q, 1)[0]
        orc.merkle_verify(root, [ol["log"]] * 4, {ol["log"]: pos}, queried, hw, cw)
        q = sorted({p >> 1 for p in q})


def test_fri_rejects_high_degree_and_bad_inputs(tb, orc):
    from tstwo_b200.channel import Blake2sChannel
    from tstwo_b200.fri import FriProver
    log, coords = secure_column_of_degree(tb, orc, 6, 1, 1)
    col = tb.SecureEvaluation(tb.CanonicCoset.new(log).circleDomain(), tb.SecureColumnByCoords.from_cpu(coords))
    with pytest.raises(ValueError, match="no columns"):
        FriProver.commit(Blake2sChannel.create(), tb.FriConfig(0, 1, 3), [])
    with pytest.raises(ValueError, match="column sizes not decreasing"):
        FriProver.commit(Blake2sChannel.create(), tb.FriConfig(0, 1, 3), [col, col])
    # a column of degree 2^6 committed as if it had degree 2^5 (blow-up 2 on a 2^7 domain): last layer must reject
    with pytest.raises(ValueError, match="invalid degree"):
        FriProver.commit(Blake2sChannel.create(), tb.FriConfig(0, 2, 3), [col])
    rnd = np.random.default_rng(3).integers(0, P, (4, 1 << 7), dtype=np.uint32)
    bad = tb.SecureEvaluation(tb.CanonicCoset.new(7).circleDomain(), tb.SecureColumnByCoords.from_cpu(list(rnd)))
    with pytest.raises(ValueError, match="invalid degree"):
        FriProver.commit(Blake2sChannel.create(), tb.FriConfig(0, 1, 3), [bad])


@pytest.mark.parametrize("with_tree", [False, True])
def test_one_call_commit_equals_layer_by_layer(tb, orc, with_tree):
    """tstwo_fri_commit_layers (transcript on the device, one host wait) against the layer-by-layer host loop: same roots, same
    layers, same channel state afterwards (incl. the queue of base-field draws), for mixed column sizes, with and without a tree."""
    from tstwo_b200.channel import Blake2sChannel
    from tstwo_b200.fri import FriProver
    cols = [secure_column_of_degree(tb, orc, d, 1, 7 * d) for d in (11, 9, 4)]
    dev_cols = [tb.SecureEvaluation(tb.CanonicCoset.new(log).circleDomain(), tb.SecureColumnByCoords.from_cpu(c)) for log, c in cols]
    tw = tb.CudaBackend.precomputeTwiddles(tb.CanonicCoset.new(cols[0][0]).half_coset()) if with_tree else None
    config = tb.FriConfig(1, 1, 4)
    res = []
    for per_layer in (True, False):
        ch = Blake2sChannel.create()
        ch.mix_u64(7)
        ch.draw_felt()                               # leaves 4 base-field draws queued: the device must continue from them
        FriProver.per_layer_calls = per_layer
        try:
            pr = FriProver.commit(ch, config, dev_cols, tw)
        finally:
            FriProver.per_layer_calls = False
        res.append((pr, ch))
    (a, ca), (b, cb) = res
    assert ca.digest() == cb.digest() and ca.getChannelTime() == cb.getChannelTime() and ca._base_queue == cb._base_queue
    assert len(a.inner_layers) == len(b.inner_layers) > 0
    for la, lb in zip(a.inner_layers, b.inner_layers):
        assert la.merkle_tree.root() == lb.merkle_tree.root()
        assert la.evaluation.domain().coset() == lb.evaluation.domain().coset()
        assert all((x == y).all() for x, y in zip(la.evaluation.values.to_cpu(), lb.evaluation.values.to_cpu()))
        for k in range(len(la.merkle_tree.layers)):
            assert (la.merkle_tree.layers[k].toCpu() == lb.merkle_tree.layers[k].toCpu()).all()
    assert a.last_layer_poly.coeffs == b.last_layer_poly.coeffs


@pytest.mark.parametrize("log_degrees,log_last,blow,n_queries", [((8, 6, 5), 2, 1, 6), ((9, 4), 0, 2, 5), ((7, 3), 2, 1, 4)])
def test_device_proof_is_accepted_by_the_host_verifier(tb, orc, log_degrees, log_last, blow, n_queries):
    """The device FriProver's proof (columns of several degrees, one joining the last layer) through oracle/fri_verify.py: the
    transcript, every layer's Merkle decommitment, the fold arithmetic and the last-layer polynomial are consistent."""
    from oracle.fri_verify import fri_verify
    from tstwo_b200.channel import Blake2sChannel
    from tstwo_b200.fri import FriProver
    cols = [secure_column_of_degree(tb, orc, d, blow, 7 * d + blow) for d in log_degrees]
    channel, vchannel = Blake2sChannel.create(), orc.Channel()
    channel.mix_u64(99); vchannel.mix_u64(99)
    dev_cols = [tb.SecureEvaluation(tb.CanonicCoset.new(log).circleDomain(), tb.SecureColumnByCoords.from_cpu(c)) for log, c in cols]
    prover = FriProver.commit(channel, tb.FriConfig(log_last, blow, n_queries), dev_cols)
    proof, by_log = prover.decommit(channel)
    layers = [(l.fri_witness, (l.decommitment.hashWitness, list(l.decommitment.columnWitness)), l.commitment)
              for l in [proof.first_layer] + proof.inner_layers]
    max_log = cols[0][0]
    evals = [[[int(c[p]) for c in col] for p in by_log[log]] for log, col in cols]
    alphas = fri_verify(orc, vchannel, log_last, blow, [log for log, _ in cols], layers, proof.last_layer_poly.coeffs,
                        by_log[max_log], evals)
    assert len(alphas) == 1 + len(proof.inner_layers)
    # the verifier's transcript ends where the prover's was before it drew the queries: same query positions
    from tstwo_b200.fri import Queries
    assert Queries.generate(vchannel, max_log, n_queries).positions == by_log[max_log]


@pytest.mark.parametrize("log_degree,n_queries", [(10, 1), (12, 9), (15, 40)])
def test_one_call_decommit_equals_layer_by_layer(tb, orc, log_degree, n_queries):
    """tstwo_fri_decommit_layers (every inner layer's positions, witnesses and Merkle walk in one call, one synchronisation) gives
    the proof the per-layer loop of FriInnerLayerProver.decommit gives; a capacity that is too small is an error."""
    import ctypes as C
    from tstwo_b200 import _lib
    from tstwo_b200.channel import Blake2sChannel
    from tstwo_b200.fri import FriProver, Queries
    log, coords = secure_column_of_degree(tb, orc, log_degree, 1, 5 * log_degree)
    col = tb.SecureEvaluation(tb.CanonicCoset.new(log).circleDomain(), tb.SecureColumnByCoords.from_cpu(coords))
    prover = FriProver.commit(Blake2sChannel.create(), tb.FriConfig(1, 1, n_queries), [col])
    rng = np.random.default_rng(n_queries)
    queries = Queries.from_positions(sorted({int(v) for v in rng.integers(0, 1 << log, n_queries)}), log)
    one = prover.decommit_on_queries(queries)
    FriProver.per_layer_calls = True
    try:
        ref = prover.decommit_on_queries(queries)
    finally:
        FriProver.per_layer_calls = False
    assert len(one.inner_layers) == len(ref.inner_layers) == log - 1 - 2
    for a, b in zip(one.inner_layers, ref.inner_layers):
        assert [list(w) for w in a.fri_witness] == [list(w) for w in b.fri_witness]
        assert list(a.decommitment.hashWitness) == list(b.decommitment.hashWitness)
        assert list(a.decommitment.columnWitness) == list(b.decommitment.columnWitness) and a.commitment == b.commitment
    # capacity check
    L = _lib.load()
    layer = prover.inner_layers[0]
    counts = (C.c_uint64 * 3)()
    w = np.zeros(4, dtype=np.uint32)
    rc = L.tstwo_fri_decommit_layers(1, _lib.dptr_array([layer.merkle_tree.layers[0].ptr]), layer.evaluation.values.ptrs(),
                                     (C.c_int * 1)(log - 1), _lib.u32_array([5]), 1, w.ctypes.data_as(_lib.u32p), 1, None, 0, None, 0, counts)
    assert rc != 0 and b"capacity" in L.tstwo_last_error()

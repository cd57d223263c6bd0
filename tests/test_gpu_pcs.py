"""prove_values and its parts (tstwo_b200/pcs.py): device grind vs the oracle's sequential search, compute_fri_quotients vs
the oracle and Stwo's `test_quotients_are_low_degree` (Rust text in core/src/pcs/quotients.ts), and the whole flow --
sample out of domain, quotients, FRI commit (whose last-layer degree check only passes if every step before it is
right), proof of work, decommit -- with every Merkle decommitment checked by the oracle's verifier."""
import hashlib

import numpy as np
import pytest

import pyfield as pf

pytestmark = pytest.mark.gpu
P = 2147483647
SECURE_GEN = ((1, 0, 478637715, 513582971), (992285211, 649143431, 740191619, 1186584352))   # circle.ts:143-146


@pytest.fixture(scope="module")
def tb():
    import tstwo_b200
    return tstwo_b200


def secure_point_double(p):
    """CirclePoint.double over QM31: (2x^2 - 1, 2xy)."""
    x, y = p
    xx = pf.qm_mul(x, x)
    return (tuple(pf.qm_sub(pf.qm_scale(xx, 2), (1, 0, 0, 0))), tuple(pf.qm_scale(pf.qm_mul(x, y), 2)))


def tz128(d: bytes) -> int:
    v = int.from_bytes(d[:16], "little")
    return 128 if v == 0 else (v & -v).bit_length() - 1


@pytest.mark.parametrize("seed,bits", [(0, 0), (1, 3), (2, 8), (3, 12), (4, 15), (5, 15)])
def test_grind_matches_sequential_search(tb, orc, seed, bits):
    from tstwo_b200.channel import Blake2sChannel
    ch, och = Blake2sChannel.create(), orc.Channel()
    ch.mix_u64(seed); och.mix_u64(seed)
    nonce = tb.grind(ch, bits)
    assert nonce == orc.grind(och, bits)
    c2 = ch.clone(); c2.mix_u64(nonce)
    assert c2.trailing_zeros() >= bits


def test_grind_many_bits(tb):
    from tstwo_b200.channel import Blake2sChannel
    ch = Blake2sChannel.create(); ch.mix_u32s([7, 7, 7])
    nonce = tb.grind(ch, 26)                      # ~2^26 compressions: several batches of the kernel
    digest = hashlib.blake2s(ch.digest() + nonce.to_bytes(8, "little")).digest()
    assert tz128(digest) >= 26
    # minimal within the last 2^16 candidates (full minimality is what the oracle comparison above pins)
    for n in range(max(0, nonce - (1 << 16)), nonce):
        assert tz128(hashlib.blake2s(ch.digest() + n.to_bytes(8, "little")).digest()) < 26
    with pytest.raises(ValueError):
        tb.grind(ch, 65)


def test_quotients_are_low_degree(tb, orc):
    LOG_SIZE, BLOW = 7, 1
    tw = tb.CudaBackend.precomputeTwiddles(tb.CanonicCoset.new(LOG_SIZE + BLOW).half_coset())
    poly = tb.CudaCirclePoly.new(np.arange(1 << LOG_SIZE, dtype=np.uint32))
    dom = tb.CanonicCoset.new(LOG_SIZE + BLOW).circleDomain()
    ev = poly.evaluateWithTwiddles(dom, tw)
    value = tb.CudaCirclePoly.eval_at_point(poly, SECURE_GEN)
    coeff = [1, 2, 3, 4]
    quot = tb.compute_fri_quotients([ev], [[tb.PointSample(SECURE_GEN, value)]], coeff, BLOW)
    assert len(quot) == 1
    # the oracle computes the same column
    init, _ = orc.canonic_half_coset(LOG_SIZE + BLOW)
    exp = orc.accumulate_quotients(LOG_SIZE + BLOW, init, [ev.values.toCpu()], coeff, [(SECURE_GEN, [(0, value)])], ts_denominators=False)
    assert all((a == b).all() for a, b in zip(quot[0].values.to_cpu(), exp))
    # each coordinate interpolates to a polynomial in the FRI space of degree 2^LOG_SIZE (is_in_fri_space)
    for c in quot[0].values.columns:
        coeffs = tb.CudaCircleEvaluation.new(dom, c.toCpu()).interpolateWithTwiddles(tw).coeffs.toCpu()
        assert not coeffs[1 << LOG_SIZE:].any() and coeffs[:1 << LOG_SIZE].any()
    # the TS CpuBackend's denominators (M31 parts only, quotients.ts:160-178) are reproduced bit for bit -- and are not low degree
    ts = tb.compute_fri_quotients([ev], [[tb.PointSample(SECURE_GEN, value)]], coeff, BLOW, ts_denominators=True)
    exp_ts = orc.accumulate_quotients(LOG_SIZE + BLOW, init, [ev.values.toCpu()], coeff, [(SECURE_GEN, [(0, value)])])
    assert all((a == b).all() for a, b in zip(ts[0].values.to_cpu(), exp_ts))
    c = tb.CudaCircleEvaluation.new(dom, ts[0].values.columns[0].toCpu()).interpolateWithTwiddles(tw).coeffs.toCpu()
    assert c[1 << LOG_SIZE:].any()
    # a wrong sampled value gives a quotient that is NOT low degree
    bad = tb.compute_fri_quotients([ev], [[tb.PointSample(SECURE_GEN, [(value[0] + 1) % P] + value[1:])]], coeff, BLOW)
    c = tb.CudaCircleEvaluation.new(dom, bad[0].values.columns[0].toCpu()).interpolateWithTwiddles(tw).coeffs.toCpu()
    assert c[1 << LOG_SIZE:].any()


def test_prove_values_flow(tb, orc):
    from tstwo_b200.channel import Blake2sChannel
    rng = np.random.default_rng(21)
    blow, pow_bits, n_queries = 1, 10, 6
    tw = tb.CudaBackend.precomputeTwiddles(tb.CanonicCoset.new(9 + blow).half_coset())
    channel = Blake2sChannel.create()
    config = tb.PcsConfig(pow_bits, tb.FriConfig(2, blow, n_queries))
    scheme = tb.CommitmentSchemeProver(config, tw)
    trace9 = [rng.integers(0, P, 1 << 9, dtype=np.uint32) for _ in range(5)]
    trace6 = [rng.integers(0, P, 1 << 6, dtype=np.uint32) for _ in range(3)]
    b = scheme.tree_builder()
    b.extend_evals([tb.CudaCircleEvaluation.new(tb.CanonicCoset.new(9).circleDomain(), c) for c in trace9])
    b.commit(channel)
    b = scheme.tree_builder()
    b.extend_evals([tb.CudaCircleEvaluation.new(tb.CanonicCoset.new(6).circleDomain(), c) for c in trace6[:2]])
    b.extend_polys([tb.CudaCirclePoly.new(trace6[2])])
    b.commit(channel)
    # out-of-domain points: the secure-field circle generator doubled a few times; a mask column is also sampled at a second point
    p0 = secure_point_double(SECURE_GEN)
    p1 = secure_point_double(p0)
    sampled_points = [[[p0]] * 5, [[p0], [p0, p1], [p1]]]
    before = channel.clone()
    proof = scheme.prove_values(sampled_points, channel)

    assert proof.commitments == scheme.roots() and len(proof.commitments) == 2
    assert [len(t) for t in proof.sampled_values] == [5, 3] and len(proof.sampled_values[1][1]) == 2
    # sampled values are the polynomials at the points (independent model: circle basis over QM31 in Python ints)
    coeffs = scheme.polynomials()[1][2].coeffs.toCpu()
    assert proof.sampled_values[1][2][0] == tb.CudaCirclePoly.eval_at_point(scheme.polynomials()[1][2], p1)
    # replay the transcript: sampled values, random coeff, FRI layers, proof of work, queries
    t = before
    t.mix_felts([v for tree in proof.sampled_values for col in tree for v in col])
    t.draw_felt()
    t.updateDigest(hashlib.blake2s(t.digest() + proof.fri_proof.first_layer.commitment).digest())
    t.draw_felt()
    for layer in proof.fri_proof.inner_layers:
        t.updateDigest(hashlib.blake2s(t.digest() + layer.commitment).digest())
        t.draw_felt()
    t.mix_felts(proof.fri_proof.last_layer_poly.coeffs)
    g = t.clone(); g.mix_u64(proof.proof_of_work)
    assert g.trailing_zeros() >= pow_bits and proof.proof_of_work == orc_grind(orc, t, pow_bits)
    t.mix_u64(proof.proof_of_work)
    from tstwo_b200.fri import Queries
    q = Queries.generate(t, 9 + blow, n_queries)
    assert t.digest() == channel.digest() and t.getChannelTime() == channel.getChannelTime()
    assert len(proof.fri_proof.last_layer_poly) == 4 and len(proof.fri_proof.inner_layers) == 9 + blow - 1 - (2 + blow)
    # the trees' decommitments verify at the FRI query positions of their column sizes
    by_log = {9 + blow: q.positions, 6 + blow: q.fold(3).positions}
    for tree_i, logs in ((0, [9 + blow] * 5), (1, [6 + blow] * 3)):
        dec = proof.decommitments[tree_i]
        orc.merkle_verify(proof.commitments[tree_i], logs, by_log, list(proof.queried_values[tree_i]), dec.hashWitness, list(dec.columnWitness))
    # queried values are the LDE values at those positions
    ev = scheme.evaluations()[0][3].values.toCpu()
    assert list(proof.queried_values[0][3::5]) == [int(ev[p]) for p in q.positions]


def orc_grind(orc, channel, bits):
    c = orc.Channel(); c.digest = channel.digest()
    return orc.grind(c, bits)

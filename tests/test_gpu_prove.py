"""GPU: prove() (tstwo_b200/prover.py) for the wide-Fibonacci component against the oracle's restatement of the same flow
(oracle/stark.py): every commitment, sampled value, FRI layer, the proof-of-work nonce and every decommitment must be equal."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
P = 2147483647


@pytest.fixture(scope="module")
def tb():
    import tstwo_b200
    return tstwo_b200


def run_prove(tb, trace, log, n_cols, pow_bits, last, blow, n_queries):
    from tstwo_b200 import air, prover
    from tstwo_b200.channel import Blake2sChannel
    tw = tb.CudaBackend.precomputeTwiddles(tb.CanonicCoset.new(log + 1 + blow).half_coset())
    channel = Blake2sChannel.create()
    scheme = tb.CommitmentSchemeProver(tb.PcsConfig(pow_bits, tb.FriConfig(last, blow, n_queries)), tw)
    b = scheme.tree_builder(); b.extend_evals([]); b.commit(channel)                      # preprocessed trace: empty
    b = scheme.tree_builder()
    dom = tb.CanonicCoset.new(log).circleDomain()
    b.extend_evals([tb.CudaCircleEvaluation.new(dom, c) for c in trace]); b.commit(channel)
    comp = air.WideFibonacciComponent(air.WideFibonacciEval.create(log, n_cols))
    return prover.prove([comp], channel, scheme), channel, scheme


# the last two are BASELINE config 5's shape (100 columns = Stwo's FIB_SEQUENCE_LENGTH, blow-up 2, 70 queries, 10 proof-of-work
# bits) at log 16 and log 20; the log-23 size itself is covered by test_config5_log23_commitments below
@pytest.mark.parametrize("log,n_cols,pow_bits,last,blow,n_queries", [(5, 6, 4, 1, 1, 3), (4, 4, 3, 0, 2, 2), (9, 20, 8, 2, 1, 5),
                                                                     (16, 100, 10, 0, 1, 70), (20, 100, 10, 0, 1, 70)])
def test_prove_matches_oracle_flow(tb, orc, log, n_cols, pow_bits, last, blow, n_queries):
    from oracle import stark
    exp = stark.prove_wide_fibonacci(log, n_cols, pow_bits, last, blow, n_queries)
    proof, channel, scheme = run_prove(tb, stark.fib_trace(log, n_cols), log, n_cols, pow_bits, last, blow, n_queries)
    p = proof.get()
    assert p.commitments == exp["commitments"]
    assert p.sampled_values == exp["sampled_values"]
    assert proof.extractCompositionOodsEval() == exp["composition_at_oods"] == exp["constraints_at_oods"]
    for got, want in zip(scheme.polynomials()[2], exp["composition_coeffs"]):
        assert (got.coeffs.toCpu() == want).all()
    assert p.proof_of_work == exp["proof_of_work"]
    assert channel.digest() == exp["channel_digest"]
    # FRI: first layer, inner layers, last layer polynomial
    layers = [p.fri_proof.first_layer] + list(p.fri_proof.inner_layers)
    assert len(layers) == len(exp["fri_layers"])
    for got, (wit, (queried, hw, cw), root) in zip(layers, exp["fri_layers"]):
        assert got.commitment == root
        assert [list(w) for w in got.fri_witness] == [list(w) for w in wit]
        assert list(got.decommitment.hashWitness) == list(hw) and list(got.decommitment.columnWitness) == list(cw)
    assert [list(c) for c in p.fri_proof.last_layer_poly.coeffs] == [list(c) for c in exp["fri"]["last_poly"]]
    # tree decommitments
    for dec, qv, (queried, hw, cw) in zip(p.decommitments, p.queried_values, exp["decommitments"]):
        assert list(qv) == list(queried) and list(dec.hashWitness) == list(hw) and list(dec.columnWitness) == list(cw)
    # the proof as BYTES: the device prover's serialisation against the oracle's own serialisation of the oracle's proof, a round
    # trip through the parser, and the reference's size estimates (prover/index.ts:306-356, 487-540)
    from tstwo_b200.prover import StarkProof
    blob = proof.toBytes()
    assert blob == stark.serialize_proof(exp, pow_bits, last, blow, n_queries), "proof bytes differ from the oracle's"
    again = StarkProof.fromBytes(blob)
    assert again.toBytes() == blob and again.commitments == p.commitments
    hashes = len(p.commitments) + sum(len(d.hashWitness) for d in p.decommitments) + \
        sum(len(l.decommitment.hashWitness) + 1 for l in [p.fri_proof.first_layer] + list(p.fri_proof.inner_layers))
    words = sum(len(q) for q in p.queried_values) + sum(len(d.columnWitness) for d in p.decommitments) + \
        sum(len(l.decommitment.columnWitness) for l in [p.fri_proof.first_layer] + list(p.fri_proof.inner_layers))
    felts = sum(len(c) for t in p.sampled_values for c in t) + len(p.fri_proof.last_layer_poly.coeffs) + \
        sum(len(l.fri_witness) for l in [p.fri_proof.first_layer] + list(p.fri_proof.inner_layers))
    assert proof.sizeEstimate() == 32 * hashes + 4 * words + 16 * felts + 8 + 64
    bd = proof.sizeBreakdownEstimate()
    assert set(bd) == {"oodsS", "queriesValues", "friSamples", "friDecommitments", "traceDecommitments"}
    assert sum(bd.values()) == proof.sizeEstimate() - 8 - 64


def test_prove_rejects_invalid_trace(tb):
    from oracle import stark
    from tstwo_b200 import prover
    bad = stark.fib_trace(5, 6)
    bad[4][7] = (int(bad[4][7]) + 1) % P
    with pytest.raises(prover.ConstraintsNotSatisfied):
        run_prove(tb, bad, 5, 6, 4, 1, 1, 3)


def test_prove_argument_checks(tb):
    from tstwo_b200 import prover
    from tstwo_b200.channel import Blake2sChannel
    tw = tb.CudaBackend.precomputeTwiddles(tb.CanonicCoset.new(6).half_coset())
    scheme = tb.CommitmentSchemeProver(tb.PcsConfig(), tw)
    with pytest.raises(TypeError, match="components must be an array"):
        prover.prove(None, Blake2sChannel.create(), scheme)
    with pytest.raises(ValueError, match="components array cannot be empty"):
        prover.prove([], Blake2sChannel.create(), scheme)
    with pytest.raises(TypeError, match="channel is required"):
        prover.prove([object()], None, scheme)
    with pytest.raises(TypeError, match="commitmentScheme is required"):
        prover.prove([object()], Blake2sChannel.create(), None)
    with pytest.raises(ValueError, match="preprocessed trace not found"):
        prover.prove([object()], Blake2sChannel.create(), scheme)


def test_config5_log23_commitments(tb, orc):
    """BASELINE config 5 at its own size: wide-Fibonacci prove() with 100 columns of 2^23 rows, blow-up 2.  The proof's trace
    commitment (and the empty preprocessed tree's) must equal the oracle's LDE + Merkle root of the same trace, the prover's own
    sanity check (composition polynomial at the out-of-domain point == constraints at the sampled values) must hold, and the
    sampled trace values must equal the oracle's eval_at_point of the oracle's coefficients for the first / last column."""
    import torch
    from oracle import stark
    from tstwo_b200 import air, prover
    from tstwo_b200.channel import Blake2sChannel
    log, n_cols, blow = 23, 100, 1
    n = 1 << log
    trace = torch.empty((n_cols, n), dtype=torch.int32, device="cuda")
    a = torch.ones(n, dtype=torch.int64, device="cuda")
    b = torch.arange(n, dtype=torch.int64, device="cuda")
    trace[0], trace[1] = a.to(torch.int32), b.to(torch.int32)
    for k in range(2, n_cols):                                   # generateTestTrace (examples/fibonacci.ts:245-258)
        a, b = b, (a * a % P + b * b % P) % P
        trace[k] = b.to(torch.int32)
    del a, b
    host = trace.cpu().numpy().view(np.uint32)
    assert (host[:, :64] == stark.fib_trace(6, n_cols)[:, :64]).all()            # same generator as the oracle's, row for row
    tw = tb.CudaBackend.precomputeTwiddles(tb.CanonicCoset.new(log + 1 + blow).half_coset())
    dom = tb.CanonicCoset.new(log).circleDomain()
    cols = [tb.CudaCircleEvaluation(dom, tb.CudaColumn.view(trace.data_ptr() + 4 * (c << log), n, trace)) for c in range(n_cols)]
    channel = Blake2sChannel.create()
    scheme = tb.CommitmentSchemeProver(tb.PcsConfig(10, tb.FriConfig(0, blow, 70)), tw)
    bld = scheme.tree_builder(); bld.extend_evals([]); bld.commit(channel)
    bld = scheme.tree_builder(); bld.extend_evals(cols); bld.commit(channel)
    proof = prover.prove([air.WideFibonacciComponent(air.WideFibonacciEval.create(log, n_cols))], channel, scheme)
    p = proof.get()
    # oracle: the same LDE + commitment (Stwo-correct transforms, as the commitment scheme uses them)
    hi, hl = orc.canonic_half_coset(log + blow)
    otw, oitw = orc.precompute_twiddles(hi, hl)
    orc.use_all_cores()
    coeffs, _, layers = orc.lde_commit(host, blow, otw, oitw, want_coeffs=True, quirks=False)
    assert p.commitments[0] == orc.merkle_commit([])[0][0].tobytes()
    assert p.commitments[1] == layers[0][0].tobytes()
    # out-of-domain samples of the first and last trace column against the oracle's coefficients at the proof's own point:
    # replay the transcript up to the point on the oracle's channel
    ch = orc.Channel()
    for root in p.commitments[:2]:
        ch.mix_root(root)
    ch.draw_felt()                                               # the composition polynomial's random coefficient
    ch.mix_root(p.commitments[2])
    oods = stark.random_point(ch)
    for c in (0, n_cols - 1):
        assert [int(v) for v in orc.eval_at_point(coeffs[c], oods[0], oods[1])] == list(p.sampled_values[1][c][0])
    assert len(p.fri_proof.inner_layers) == log + 1 + blow - 1 - blow and len(p.sampled_values[2]) == 4


@pytest.mark.parametrize("log,n_cols,pow_bits,last,blow,n_queries", [(5, 6, 4, 1, 1, 3), (6, 9, 6, 0, 2, 7), (12, 32, 8, 1, 1, 20), (16, 100, 10, 0, 1, 70)])
def test_device_proof_bytes_are_accepted_by_the_whole_proof_verifier(tb, orc, log, n_cols, pow_bits, last, blow, n_queries):
    """The device prover's proof, as bytes, through oracle/stark_verify.py: transcript, out-of-domain check, every Merkle
    decommitment, DEEP quotients at the queries, FRI down to the last layer, proof of work.  A flipped byte is rejected."""
    from oracle import stark
    from oracle.stark_verify import verify_wide_fibonacci
    proof, _, _ = run_prove(tb, stark.fib_trace(log, n_cols), log, n_cols, pow_bits, last, blow, n_queries)
    blob = proof.toBytes()
    verify_wide_fibonacci(blob, log, n_cols)
    bad = bytearray(blob); bad[len(bad) // 2] ^= 4
    with pytest.raises(ValueError):
        verify_wide_fibonacci(bytes(bad), log, n_cols)

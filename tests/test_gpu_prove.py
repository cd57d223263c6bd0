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


@pytest.mark.parametrize("log,n_cols,pow_bits,last,blow,n_queries", [(5, 6, 4, 1, 1, 3), (4, 4, 3, 0, 2, 2), (9, 20, 8, 2, 1, 5)])
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

"""The proof byte format (tstwo_b200/proof_format.py; the reference defines none) and the reference's size estimates
(prover/index.ts:185-372, 471-540) on the host: the oracle's proof serialised by the oracle's own writer parses with the product's
reader and re-serialises to the same bytes; damaged inputs are rejected."""
import pytest


@pytest.mark.parametrize("log,n_cols,pow_bits,last,blow,n_queries", [(5, 6, 4, 1, 1, 3), (4, 4, 3, 0, 2, 2), (7, 10, 5, 2, 1, 6)])
def test_oracle_proof_bytes_round_trip(log, n_cols, pow_bits, last, blow, n_queries):
    from oracle import stark
    from tstwo_b200.prover import StarkProof
    exp = stark.prove_wide_fibonacci(log, n_cols, pow_bits, last, blow, n_queries)
    blob = stark.serialize_proof(exp, pow_bits, last, blow, n_queries)
    proof = StarkProof.fromBytes(blob)
    p = proof.get()
    assert proof.toBytes() == blob
    assert p.commitments == exp["commitments"] and p.proof_of_work == exp["proof_of_work"]
    assert p.sampled_values == exp["sampled_values"]
    assert (p.config.pow_bits, p.config.fri_config.log_last_layer_degree_bound, p.config.fri_config.log_blowup_factor,
            p.config.fri_config.n_queries) == (pow_bits, last, blow, n_queries)
    assert [list(c) for c in p.fri_proof.last_layer_poly.coeffs] == [list(c) for c in exp["fri"]["last_poly"]]
    assert proof.extractCompositionOodsEval() == exp["composition_at_oods"]
    # size estimates: every variable-length part is counted once, plus 8 bytes of nonce and 64 of config
    bd = proof.sizeBreakdownEstimate()
    assert sum(bd.values()) + 8 + 64 == proof.sizeEstimate()
    assert bd["oodsS"] == 16 * sum(len(c) for t in p.sampled_values for c in t)
    assert bd["queriesValues"] == 4 * sum(len(q) for q in p.queried_values)
    # the byte string is the estimate plus framing (magic, config words, length prefixes), minus the 64-byte config estimate
    assert len(blob) > proof.sizeEstimate() - 64


def test_damaged_proof_bytes_are_rejected():
    from oracle import stark
    from tstwo_b200.prover import StarkProof
    blob = stark.serialize_proof(stark.prove_wide_fibonacci(4, 4, 3, 0, 2, 2), 3, 0, 2, 2)
    with pytest.raises(ValueError, match="bad magic"):
        StarkProof.fromBytes(b"X" + blob[1:])
    with pytest.raises(ValueError, match="truncated"):
        StarkProof.fromBytes(blob[:-5])
    with pytest.raises(ValueError, match="trailing"):
        StarkProof.fromBytes(blob + b"\x00")

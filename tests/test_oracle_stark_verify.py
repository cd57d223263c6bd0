"""A whole-proof verifier (oracle/stark_verify.py: transcript, out-of-domain check, Merkle decommitments, DEEP quotients at the
queries -> FRI, proof of work) accepts the oracle prover's proofs, taken as BYTES, and rejects tampered ones."""
import struct

import pytest


def _proof(log, n_cols, pow_bits, last, blow, n_queries):
    from oracle import stark
    exp = stark.prove_wide_fibonacci(log, n_cols, pow_bits, last, blow, n_queries)
    return exp, stark.serialize_proof(exp, pow_bits, last, blow, n_queries)


@pytest.mark.parametrize("log,n_cols,pow_bits,last,blow,n_queries", [(5, 6, 4, 1, 1, 3), (4, 4, 3, 0, 2, 2), (7, 10, 5, 2, 1, 6), (9, 20, 8, 0, 1, 12)])
def test_oracle_proofs_verify(log, n_cols, pow_bits, last, blow, n_queries):
    from oracle.stark_verify import verify_wide_fibonacci
    _, blob = _proof(log, n_cols, pow_bits, last, blow, n_queries)
    verify_wide_fibonacci(blob, log, n_cols)


def test_tampered_proofs_are_rejected():
    from oracle.stark_verify import parse_proof, verify_wide_fibonacci
    log, n_cols = 6, 8
    exp, blob = _proof(log, n_cols, 5, 1, 1, 5)
    verify_wide_fibonacci(blob, log, n_cols)
    pr = parse_proof(blob)
    # byte offsets of a few fields, walked like the format: magic 8 + config 16 + commitments (4 + 3 * 32)
    off_sampled = 8 + 16 + 4 + 96
    # first sampled value of the trace tree: trees: u32 n_trees, tree 0: u32 0 columns; tree 1: u32 n_cols, column 0: u32 1, 16 bytes
    off_first_sample = off_sampled + 4 + 4 + 4 + 4
    def flip(data, off):
        b = bytearray(data); b[off] ^= 1
        return bytes(b)
    with pytest.raises(ValueError, match="OodsNotMatching"):
        verify_wide_fibonacci(flip(blob, off_first_sample), log, n_cols)
    # a commitment: the transcript changes, so the out-of-domain point does, and the samples no longer fit
    with pytest.raises(ValueError):
        verify_wide_fibonacci(flip(blob, 8 + 16 + 4 + 40), log, n_cols)
    # the last byte belongs to the last-layer polynomial
    with pytest.raises(ValueError):
        verify_wide_fibonacci(flip(blob, len(blob) - 1), log, n_cols)
    # a queried value of the trace tree (located by re-serialising the parsed proof up to it)
    marker = struct.pack("<I", len(pr["queried"][1])) + struct.pack(f"<{len(pr['queried'][1])}I", *pr["queried"][1])
    at = blob.index(marker)
    with pytest.raises(ValueError, match="RootMismatch"):
        verify_wide_fibonacci(flip(blob, at + 4 + 8), log, n_cols)
    # another nonce: with 5 bits most nonces fail the proof of work; the rest change the queries
    nonce_bytes = struct.pack("<Q", pr["nonce"])
    real = blob.rindex(nonce_bytes, 0, len(blob))
    bad = blob[:real] + struct.pack("<Q", pr["nonce"] + 1) + blob[real + 8:]
    with pytest.raises(ValueError):
        verify_wide_fibonacci(bad, log, n_cols)
    # wrong statement
    with pytest.raises(ValueError):
        verify_wide_fibonacci(blob, log, n_cols + 1)
    with pytest.raises(ValueError, match="BadMagic"):
        verify_wide_fibonacci(b"Y" + blob[1:], log, n_cols)

"""TEST INFRASTRUCTURE (oracle): a whole-proof verifier for the wide-Fibonacci STARK, so that the proofs of the device prover and
of the oracle's prover (oracle/stark.py) are checked for what a proof is for.  The reference's verifier is a skeleton
(prover/index.ts verify() and pcs/verifier.ts are placeholders around the Rust text); this follows Stwo's flow as that text
gives it (pcs/verifier.ts Rust comment, prover/index.ts:582-666 mirrored):

  transcript   mix_root(preprocessed), mix_root(trace); random_coeff = draw_felt; mix_root(composition); oods point;
               mix_felts(sampled values); fri_random_coeff = draw_felt; FRI commit phase; proof of work; queries
  checks       composition polynomial at the oods point == constraints at the sampled mask values;
               every tree's Merkle decommitment against its commitment (oracle.merkle_verify);
               DEEP quotients at the query positions from the QUERIED values -> FRI (oracle/fri_verify.py) down to the last layer;
               trailing zeros of the proof-of-work digest.

Input: the proof as BYTES in the format of tstwo_b200/proof_format.py, parsed here independently of the product's reader."""
from __future__ import annotations

import struct

import numpy as np

import oracle as orc
from oracle import stark
from oracle.fri_verify import fri_verify

P = 2147483647


class _R:
    def __init__(self, data: bytes):
        self.d, self.o = data, 0

    def take(self, n):
        if self.o + n > len(self.d):
            raise ValueError("ProofTruncated")
        b = self.d[self.o:self.o + n]; self.o += n
        return b

    def u32(self): return struct.unpack("<I", self.take(4))[0]
    def words(self, n): return list(struct.unpack(f"<{n}I", self.take(4 * n)))


def parse_proof(data: bytes) -> dict:
    r = _R(data)
    if r.take(8) != b"TSTWOPF1":
        raise ValueError("BadMagic")
    pow_bits, last, blow, n_queries = struct.unpack("<IIII", r.take(16))
    commitments = [r.take(32) for _ in range(r.u32())]
    sampled = []
    for _ in range(r.u32()):
        sampled.append([[r.words(4) for _ in range(r.u32())] for _ in range(r.u32())])
    def decommitment():
        hw = [r.take(32) for _ in range(r.u32())]
        return hw, r.words(r.u32())
    decommitments = [decommitment() for _ in range(r.u32())]
    queried = [r.words(r.u32()) for _ in range(r.u32())]
    nonce = struct.unpack("<Q", r.take(8))[0]
    def fri_layer():
        w = [r.words(4) for _ in range(r.u32())]
        d = decommitment()
        return w, d, r.take(32)
    first = fri_layer()
    inner = [fri_layer() for _ in range(r.u32())]
    last_poly = [r.words(4) for _ in range(r.u32())]
    if r.o != len(r.d):
        raise ValueError("TrailingBytes")
    return {"pow_bits": pow_bits, "last": last, "blow": blow, "n_queries": n_queries, "commitments": commitments, "sampled": sampled,
            "decommitments": decommitments, "queried": queried, "nonce": nonce, "fri_layers": [first] + inner, "last_poly": last_poly}


def verify_wide_fibonacci(proof_bytes: bytes, log_n_rows: int, n_cols: int) -> None:
    """Raises ValueError naming the failed check; returns None when the proof is accepted."""
    pr = parse_proof(proof_bytes)
    blow = pr["blow"]
    comp_log = log_n_rows + 1
    trace_log, top_log = log_n_rows + blow, comp_log + blow
    tree_logs = [[], [trace_log] * n_cols, [top_log] * 4]
    if len(pr["commitments"]) != 3 or [len(t) for t in pr["sampled"]] != [0, n_cols, 4] or any(len(c) != 1 for t in pr["sampled"] for c in t):
        raise ValueError("InvalidProofStructure")
    channel = orc.Channel()
    channel.mix_root(pr["commitments"][0])
    channel.mix_root(pr["commitments"][1])
    random_coeff = channel.draw_felt()
    channel.mix_root(pr["commitments"][2])
    oods = stark.random_point(channel)

    # composition polynomial at the point (from its four coordinate samples) == the constraints at the sampled trace values
    qm = stark._qm
    comp = pr["sampled"][2][0][0]
    for k, unit in ((1, [0, 1, 0, 0]), (2, [0, 0, 1, 0]), (3, [0, 0, 0, 1])):
        comp = qm(0, comp, qm(2, pr["sampled"][2][k][0], unit))
    denom_inv = qm(4, stark.coset_vanishing_secure(orc.subgroup_gen(log_n_rows + 1), log_n_rows, oods))
    acc = [0, 0, 0, 0]
    for i in range(n_cols - 2):
        a, b, c = pr["sampled"][1][i][0], pr["sampled"][1][i + 1][0], pr["sampled"][1][i + 2][0]
        acc = qm(0, qm(2, acc, random_coeff), qm(2, denom_inv, qm(1, c, qm(0, qm(2, a, a), qm(2, b, b)))))
    if [int(v) for v in comp] != [int(v) for v in acc]:
        raise ValueError("OodsNotMatching")

    channel.mix_felts([v for tree in pr["sampled"] for col in tree for v in col])
    fri_random_coeff = channel.draw_felt()
    half = lambda log: orc.canonic_half_coset(log)[0]
    fold = lambda pos, k: sorted({p >> k for p in pos})

    def after_fri_commit(ch):
        # proof of work, then the queries (pcs/prover.ts:85-156 Rust text order)
        probe = orc.Channel(); probe.digest = ch.digest
        probe.mix_u64(pr["nonce"])
        if probe.trailing_zeros() < pr["pow_bits"]:
            raise ValueError("ProofOfWork")
        ch.mix_u64(pr["nonce"])
        positions = stark.generate_queries(ch, top_log, pr["n_queries"])
        by_log = {log: fold(positions, top_log - log) for log in {trace_log, top_log}}
        # every commitment tree opens at those positions
        for ti, logs in enumerate(tree_logs):
            hw, cw = pr["decommitments"][ti]
            if logs:
                orc.merkle_verify(pr["commitments"][ti], logs, by_log, pr["queried"][ti], hw, cw)
            elif hw or cw or pr["queried"][ti]:
                raise ValueError("WitnessTooLong")
        # DEEP quotients at the query positions from the queried values: the quotient of a row depends on that row only, so the
        # column-wise oracle routine is run on columns that hold the queried values at the queried rows (zeros elsewhere)
        evals = []
        for log, ti in ((top_log, 2), (trace_log, 1)):
            pos, k = by_log[log], len(tree_logs[ti])
            if len(pr["queried"][ti]) != k * len(pos):
                raise ValueError("TooFewQueriedValues")
            cols = [np.zeros(1 << log, dtype=np.uint32) for _ in range(k)]
            for j, q in enumerate(pos):
                for c in range(k):
                    cols[c][q] = pr["queried"][ti][j * k + c]
            batches = [(oods, [(c, pr["sampled"][ti][c][0]) for c in range(k)])]
            quot = orc.accumulate_quotients(log, half(log), cols, fri_random_coeff, batches, ts_denominators=False)
            evals.append([[int(quot[m][q]) for m in range(4)] for q in pos])
        return positions, evals

    column_logs = [top_log, trace_log] if top_log != trace_log else [top_log]
    layers = [(w, d, root) for w, d, root in pr["fri_layers"]]
    fri_verify(orc, channel, pr["last"], blow, column_logs, layers, pr["last_poly"], after_fri_commit, None)

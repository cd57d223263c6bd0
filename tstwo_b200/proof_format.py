"""Proof size estimates and a byte format for StarkProof.

Size estimates follow the reference: `ProverStarkProof.sizeEstimate` / `sizeBreakdownEstimate` and the `sizeEstimate*` helpers
(packages/core/src/prover/index.ts:185-372, 471-540; Stwo's `SizeEstimate` trait): a hash counts 32 bytes, a base-field element 4,
a secure-field element 16, the proof-of-work nonce 8, the config 64.

The BYTE FORMAT is defined here: the reference has none (prover/index.ts:606-619 is a stub and nothing serialises a proof), but a
prover backend whose contract is "proof bytes bit-exact" needs bytes to compare.  It is the plain little-endian walk of Stwo's
`CommitmentSchemeProof` in field order, every variable-length sequence preceded by its u32 length:

    magic "TSTWOPF1"
    config                pow_bits u32, log_last_layer_degree_bound u32, log_blowup_factor u32, n_queries u32
    commitments           u32 n, then n x 32 bytes
    sampled_values        u32 n_trees; per tree u32 n_columns; per column u32 n_samples, then n_samples x 4 x u32 (QM31, to_m31_array order)
    decommitments         u32 n_trees; per tree MerkleDecommitment = u32 n_hashes, n_hashes x 32 bytes, u32 n_words, n_words x u32
    queried_values        u32 n_trees; per tree u32 n_words, n_words x u32
    proof_of_work         u64
    fri_proof             FriLayerProof (first layer), u32 n_inner, n_inner x FriLayerProof, u32 n_coeffs, n_coeffs x 4 x u32 (last layer poly)
    FriLayerProof         u32 n_witness, n_witness x 4 x u32, MerkleDecommitment, 32 bytes commitment

`tests/test_gpu_prove.py` compares these bytes with the oracle's own serialisation of the oracle's proof (oracle/stark.py)."""
from __future__ import annotations

import struct
from typing import List

import numpy as np

MAGIC = b"TSTWOPF1"
HASH_BYTES, BASE_FIELD_BYTES, SECURE_FIELD_BYTES, POW_BYTES, CONFIG_BYTES = 32, 4, 16, 8, 64


# ---- size estimates -----------------------------------------------------------------------------------------------------------
def _n_words(words) -> int:
    return int(np.asarray(words).size)


def size_estimate_decommitment(d) -> int:
    """sizeEstimateMerkleDecommitment, prover/index.ts:232-250."""
    return HASH_BYTES * len(d.hashWitness) + BASE_FIELD_BYTES * _n_words(d.columnWitness)


def size_estimate_fri_layer(layer) -> int:
    """sizeEstimateFriLayerProof, prover/index.ts:256-277."""
    return SECURE_FIELD_BYTES * len(layer.fri_witness) + size_estimate_decommitment(layer.decommitment) + HASH_BYTES


def size_estimate_fri_proof(fri) -> int:
    """sizeEstimateFriProof, prover/index.ts:283-300."""
    return (size_estimate_fri_layer(fri.first_layer) + sum(size_estimate_fri_layer(l) for l in fri.inner_layers)
            + SECURE_FIELD_BYTES * len(fri.last_layer_poly.coeffs))


def _sampled_bytes(sampled_values) -> int:
    return SECURE_FIELD_BYTES * sum(len(col) for tree in sampled_values for col in tree)


def size_estimate(p) -> int:
    """sizeEstimateCommitmentSchemeProof, prover/index.ts:306-356."""
    return (HASH_BYTES * len(p.commitments) + _sampled_bytes(p.sampled_values) + sum(size_estimate_decommitment(d) for d in p.decommitments)
            + BASE_FIELD_BYTES * sum(_n_words(q) for q in p.queried_values) + POW_BYTES + size_estimate_fri_proof(p.fri_proof) + CONFIG_BYTES)


def size_breakdown_estimate(p) -> dict:
    """ProverStarkProof.sizeBreakdownEstimate, prover/index.ts:487-540 (same five keys)."""
    fri = p.fri_proof
    inner_samples = SECURE_FIELD_BYTES * sum(len(l.fri_witness) for l in fri.inner_layers)
    inner_hashes = sum(size_estimate_decommitment(l.decommitment) + HASH_BYTES for l in fri.inner_layers)
    return {
        "oodsS": _sampled_bytes(p.sampled_values),
        "queriesValues": BASE_FIELD_BYTES * sum(_n_words(q) for q in p.queried_values),
        "friSamples": SECURE_FIELD_BYTES * len(fri.last_layer_poly.coeffs) + inner_samples + SECURE_FIELD_BYTES * len(fri.first_layer.fri_witness),
        "friDecommitments": inner_hashes + size_estimate_decommitment(fri.first_layer.decommitment) + HASH_BYTES,
        "traceDecommitments": HASH_BYTES * len(p.commitments) + sum(size_estimate_decommitment(d) for d in p.decommitments),
    }


# ---- bytes --------------------------------------------------------------------------------------------------------------------
def _u32(v: int) -> bytes:
    return struct.pack("<I", v)


def _words(words) -> bytes:
    return np.ascontiguousarray(np.asarray(words, dtype="<u4")).tobytes()


def _hashes(hashes) -> bytes:
    rows = getattr(hashes, "_rows", None)                 # HashRows: one copy instead of one bytes object per hash
    if rows is not None:
        return np.ascontiguousarray(rows).tobytes()
    out = b"".join(bytes(h) for h in hashes)
    if len(out) != HASH_BYTES * len(hashes):
        raise ValueError("a hash is not 32 bytes")
    return out


def _decommitment(d) -> List[bytes]:
    cw = np.asarray(d.columnWitness, dtype="<u4").reshape(-1)
    return [_u32(len(d.hashWitness)), _hashes(d.hashWitness), _u32(cw.size), cw.tobytes()]


def _fri_layer(layer) -> List[bytes]:
    w = np.asarray(layer.fri_witness, dtype="<u4").reshape(-1, 4) if len(layer.fri_witness) else np.zeros((0, 4), dtype="<u4")
    commitment = bytes(layer.commitment)
    if len(commitment) != HASH_BYTES:
        raise ValueError("a commitment is not 32 bytes")
    return [_u32(w.shape[0]), w.tobytes()] + _decommitment(layer.decommitment) + [commitment]


def to_bytes(p) -> bytes:
    """The byte string of a CommitmentSchemeProof in the format of the module docstring."""
    cfg, fri_cfg = p.config, p.config.fri_config
    out: List[bytes] = [MAGIC, struct.pack("<IIII", cfg.pow_bits, fri_cfg.log_last_layer_degree_bound, fri_cfg.log_blowup_factor, fri_cfg.n_queries)]
    out.append(_u32(len(p.commitments)))
    out.append(_hashes(p.commitments))
    out.append(_u32(len(p.sampled_values)))
    for tree in p.sampled_values:
        out.append(_u32(len(tree)))
        for col in tree:
            out.append(_u32(len(col)))
            if len(col):
                out.append(_words(np.asarray(col, dtype="<u4").reshape(-1, 4)))
    out.append(_u32(len(p.decommitments)))
    for d in p.decommitments:
        out += _decommitment(d)
    out.append(_u32(len(p.queried_values)))
    for q in p.queried_values:
        qa = np.asarray(q, dtype="<u4").reshape(-1)
        out += [_u32(qa.size), qa.tobytes()]
    out.append(struct.pack("<Q", p.proof_of_work))
    fri = p.fri_proof
    out += _fri_layer(fri.first_layer)
    out.append(_u32(len(fri.inner_layers)))
    for layer in fri.inner_layers:
        out += _fri_layer(layer)
    coeffs = np.asarray(fri.last_layer_poly.coeffs, dtype="<u4").reshape(-1, 4)
    out += [_u32(coeffs.shape[0]), coeffs.tobytes()]
    return b"".join(out)


class _Reader:
    def __init__(self, data: bytes):
        self.d, self.o = memoryview(data), 0

    def take(self, n: int) -> bytes:
        if self.o + n > len(self.d):
            raise ValueError("proof bytes truncated")
        b = bytes(self.d[self.o:self.o + n]); self.o += n
        return b

    def u32(self) -> int: return struct.unpack("<I", self.take(4))[0]
    def words(self, n: int) -> np.ndarray: return np.frombuffer(self.take(4 * n), dtype="<u4").astype(np.uint32)
    def hashes(self, n: int) -> List[bytes]: return [self.take(HASH_BYTES) for _ in range(n)]


def from_bytes(data: bytes):
    """Inverse of to_bytes: a CommitmentSchemeProof with plain host values (lists, numpy arrays)."""
    from .backend import FriConfig, MerkleDecommitment, PcsConfig
    from .fri import FriLayerProof, FriProof, LinePoly
    from .pcs import CommitmentSchemeProof
    r = _Reader(data)
    if r.take(len(MAGIC)) != MAGIC:
        raise ValueError("not a tstwo proof (bad magic)")
    pow_bits, last, blow, n_queries = struct.unpack("<IIII", r.take(16))
    config = PcsConfig(pow_bits, FriConfig(last, blow, n_queries))
    commitments = r.hashes(r.u32())
    sampled = [[[[int(v) for v in r.words(4)] for _ in range(r.u32())] for _ in range(r.u32())] for _ in range(r.u32())]

    def decommitment():
        hw = r.hashes(r.u32())
        return MerkleDecommitment(hw, r.words(r.u32()))

    decommitments = [decommitment() for _ in range(r.u32())]
    queried = [r.words(r.u32()) for _ in range(r.u32())]
    pow_nonce = struct.unpack("<Q", r.take(8))[0]

    def fri_layer():
        w = [[int(v) for v in r.words(4)] for _ in range(r.u32())]
        d = decommitment()
        return FriLayerProof(w, d, r.take(HASH_BYTES))

    first = fri_layer()
    inner = [fri_layer() for _ in range(r.u32())]
    last_poly = LinePoly([[int(v) for v in r.words(4)] for _ in range(r.u32())])
    if r.o != len(r.d):
        raise ValueError("trailing bytes after the proof")
    return CommitmentSchemeProof(config, commitments, sampled, decommitments, queried, pow_nonce, FriProof(first, inner, last_poly))

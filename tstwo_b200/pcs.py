"""The rest of the commitment scheme's prover around the hot path: out-of-domain sampling, DEEP quotients, FRI, proof of
work and decommitment -- `CommitmentSchemeProver.prove_values` of Stwo, which the reference carries only as Rust text
(packages/core/src/pcs/prover.ts:85-156) next to `compute_fri_quotients` / `ColumnSampleBatch` (pcs/quotients.ts, Rust
text) and a CPU `grind` (backend/cpu/grind.ts:31-42).

Device work: eval_at_point per polynomial (tstwo_eval_at_point), accumulate_quotients per column size
(tstwo_accumulate_quotients), FriProver (fri.py), the nonce search (tstwo_grind), Merkle decommitments
(tstwo_gather_words).  Host work: the transcript and the bookkeeping of which column is sampled where."""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Sequence, Tuple

from . import _lib
from ._lib import check
from .backend import (_L, CommitmentSchemeProver, CudaBackend, CudaCircleEvaluation, CudaCirclePoly, MerkleDecommitment, PcsConfig,
                      SecureEvaluation, accumulate_quotients)
from .circle import CanonicCoset
from .fri import FriProof, FriProver

Point = Tuple[Sequence[int], Sequence[int]]        # CirclePoint<SecureField>: (x[4], y[4])


class PointSample:
    """pcs/quotients.ts (Rust text): a column's value at an out-of-domain point."""

    def __init__(self, point: Point, value: Sequence[int]):
        self.point = (tuple(int(w) for w in point[0]), tuple(int(w) for w in point[1]))
        self.value = [int(w) for w in value]


class ColumnSampleBatch:
    """All columns sampled at one point: (point, [(column index, value)])."""

    def __init__(self, point: Point, columns_and_values: List[Tuple[int, List[int]]]):
        self.point, self.columns_and_values = point, columns_and_values

    @staticmethod
    def new_vec(samples: Sequence[Sequence[PointSample]]) -> List["ColumnSampleBatch"]:
        """Groups samples by point, first appearance first (the Rust text's IndexMap keeps insertion order)."""
        grouped: Dict[Point, List[Tuple[int, List[int]]]] = {}
        for column_index, column_samples in enumerate(samples):
            for s in column_samples:
                grouped.setdefault(s.point, []).append((column_index, s.value))
        return [ColumnSampleBatch(p, cv) for p, cv in grouped.items()]


def compute_fri_quotients(columns: Sequence[CudaCircleEvaluation], samples: Sequence[Sequence[PointSample]], random_coeff: Sequence[int],
                          log_blowup_factor: int, ts_denominators: bool = False) -> List[SecureEvaluation]:
    """pcs/quotients.ts `compute_fri_quotients`: columns grouped by log size, largest first (stable), one
    accumulate_quotients launch per size on the canonic domain of that size.  Uses Stwo's denominators by default: the TS
    CpuBackend's (ts_denominators=True) do not give low-degree quotients, so FRI would reject them."""
    if len(columns) != len(samples):
        raise ValueError("one sample list per column")
    order = sorted(range(len(columns)), key=lambda i: -columns[i].domain.logSize())
    out, i = [], 0
    while i < len(order):
        log = columns[order[i]].domain.logSize()
        group = []
        while i < len(order) and columns[order[i]].domain.logSize() == log:
            group.append(order[i]); i += 1
        batches = ColumnSampleBatch.new_vec([samples[k] for k in group])
        out.append(accumulate_quotients(CanonicCoset.new(log).circleDomain(), [columns[k] for k in group], random_coeff,
                                        [(b.point, b.columns_and_values) for b in batches], log_blowup_factor, ts_denominators))
    return out


def grind(channel, pow_bits: int) -> int:
    """backend/cpu/grind.ts:31-42 on the device: smallest nonce whose mix_u64 leaves >= pow_bits trailing zero bits."""
    if not isinstance(pow_bits, int) or pow_bits < 0 or pow_bits > 64:
        raise ValueError("pow_bits must be in 0..64")
    nonce = C.c_uint64()
    digest = (C.c_uint8 * 32)(*channel.digest())
    check(_L().tstwo_grind(digest, pow_bits, C.byref(nonce)))
    return int(nonce.value)


CudaBackend.grind = staticmethod(grind)


class CommitmentSchemeProof:
    """pcs/prover.ts:158-168 (Rust text)."""

    def __init__(self, config: PcsConfig, commitments: List[bytes], sampled_values, decommitments: List[MerkleDecommitment],
                 queried_values, proof_of_work: int, fri_proof: FriProof):
        self.config, self.commitments, self.sampled_values = config, commitments, sampled_values
        self.decommitments, self.queried_values, self.proof_of_work, self.fri_proof = decommitments, queried_values, proof_of_work, fri_proof


def prove_values(self: CommitmentSchemeProver, sampled_points, channel) -> CommitmentSchemeProof:
    """pcs/prover.ts:85-156 (Rust text).  sampled_points[tree][column] = list of points (x[4], y[4])."""
    # every (polynomial, point) pair, evaluated point by point in batches (tstwo_eval_at_point_many)
    by_point: Dict[Point, List[Tuple[int, int, int]]] = {}
    for ti, (tree, tree_pts) in enumerate(zip(self.trees, sampled_points)):
        if len(tree_pts) != len(tree.polynomials):
            raise ValueError("one list of sample points per column")
        for ci, pts in enumerate(tree_pts):
            for k, pt in enumerate(pts):
                key = (tuple(int(w) for w in pt[0]), tuple(int(w) for w in pt[1]))
                by_point.setdefault(key, []).append((ti, ci, k))
    values = {}
    for pt, where in by_point.items():
        res = CudaCirclePoly.eval_many_at_point([self.trees[ti].polynomials[ci] for ti, ci, _ in where], pt)
        for w, v in zip(where, res):
            values[w] = v
    samples = [[[PointSample(pt, values[(ti, ci, k)]) for k, pt in enumerate(pts)] for ci, pts in enumerate(tree_pts)]
               for ti, tree_pts in enumerate(sampled_points)]
    sampled_values = [[[s.value for s in col] for col in tree] for tree in samples]
    channel.mix_felts([v for tree in sampled_values for col in tree for v in col])
    columns = [e for tree in self.trees for e in tree.evaluations]
    flat_samples = [col for tree in samples for col in tree]
    quotients = compute_fri_quotients(columns, flat_samples, channel.draw_felt(), self.config.fri_config.log_blowup_factor)
    fri_prover = FriProver.commit(channel, self.config.fri_config, quotients, self.twiddles)
    proof_of_work = grind(channel, self.config.pow_bits)
    channel.mix_u64(proof_of_work)
    fri_proof, query_positions_per_log_size = fri_prover.decommit(channel)
    results = [tree.decommit(query_positions_per_log_size) for tree in self.trees]
    return CommitmentSchemeProof(self.config, self.roots(), sampled_values, [d for _, d in results], [v for v, _ in results],
                                 proof_of_work, fri_proof)


CommitmentSchemeProver.prove_values = prove_values

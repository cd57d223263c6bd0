"""prove() over the device commitment scheme (prover/index.ts:582-666).

The reference's prove() has the right skeleton but simulated components ("For now, we'll simulate ...", :606-616, :632);
this module fills the placeholders in with the real calls the comments name, as Stwo's prove does:
  random_coeff = channel.draw_felt()
  composition polynomial = sum over components of their constraint quotients on the evaluation domain, finalize()
  commit its 4 coordinate polynomials as the last tree
  oods_point = CirclePoint.get_random_point(channel)
  sample trace mask points + the composition polynomial at oods_point; CommitmentSchemeProver.prove_values
  sanity check: composition value recombined from the samples == the constraints evaluated at the sampled mask values.
A component is any object with nConstraints(), maxConstraintLogDegreeBound(), maskPoints(point),
evaluateConstraintQuotientsOnDomain(trace, accumulator) and evaluateConstraintQuotientsAtPoint(point, mask, accumulator)
(air/index.ts:38-95 Component / ComponentProver); air.WideFibonacciComponent is one."""
from __future__ import annotations

from typing import List, Sequence

from .air import DomainEvaluationAccumulator, PointEvaluationAccumulator, SecureCirclePoly, from_partial_evals, get_random_point
from .backend import CommitmentSchemeProver
from .pcs import CommitmentSchemeProof

PREPROCESSED_TRACE_IDX = 0          # prover/index.ts PREPROCESSED_TRACE_IDX
SECURE_EXTENSION_DEGREE = 4


class ConstraintsNotSatisfied(Exception):
    """ProvingError::ConstraintsNotSatisfied (prover/index.ts ProverProvingErrorException.constraintsNotSatisfied)."""

    def __init__(self):
        super().__init__("Constraints not satisfied.")


class Trace:
    """air/index.ts Trace: polynomials and evaluations of every committed tree."""

    def __init__(self, polys, evals):
        self.polys, self.evals = polys, evals


class ComponentProvers:
    """air/components.ts: the components of a proof taken together."""

    def __init__(self, components: Sequence, n_preprocessed_columns: int = 0):
        self.components = list(components)
        self.nPreprocessedColumns = n_preprocessed_columns

    def compositionLogDegreeBound(self) -> int:
        return max(c.maxConstraintLogDegreeBound() for c in self.components)

    def nConstraints(self) -> int:
        return sum(c.nConstraints() for c in self.components)

    def maskPoints(self, point):
        """Per tree, the mask points of every component's columns, components in order (components.ts:51-79)."""
        trees: List[List] = []
        for c in self.components:
            for t, cols in enumerate(c.maskPoints(point)):
                while len(trees) <= t:
                    trees.append([])
                trees[t].extend(cols)
        return trees

    def evalCompositionPolynomialAtPoint(self, point, mask_values, random_coeff) -> List[int]:
        acc = PointEvaluationAccumulator.new(random_coeff)
        for c in self.components:
            c.evaluateConstraintQuotientsAtPoint(point, mask_values, acc)
        return acc.finalize()

    def computeCompositionPolynomial(self, random_coeff, trace: Trace, twiddles=None) -> SecureCirclePoly:
        acc = DomainEvaluationAccumulator.new(random_coeff, self.compositionLogDegreeBound(), self.nConstraints())
        for c in self.components:
            c.evaluateConstraintQuotientsOnDomain(trace, acc, twiddles)
        return acc.finalize()


class StarkProof:
    """prover/index.ts:480-566: wraps the commitment-scheme proof."""

    def __init__(self, commitment_scheme_proof: CommitmentSchemeProof):
        self.commitmentSchemeProof = commitment_scheme_proof

    create = staticmethod(lambda p: StarkProof(p))

    def get(self) -> CommitmentSchemeProof: return self.commitmentSchemeProof

    @property
    def sampledValues(self): return self.commitmentSchemeProof.sampled_values

    @property
    def commitments(self): return self.commitmentSchemeProof.commitments

    def sizeEstimate(self) -> int:
        """prover/index.ts:471-476."""
        from . import proof_format
        if getattr(self, "_size_estimate_cache", None) is None:
            self._size_estimate_cache = proof_format.size_estimate(self.commitmentSchemeProof)
        return self._size_estimate_cache
    size_estimate = sizeEstimate

    def sizeBreakdownEstimate(self) -> dict:
        """prover/index.ts:487-540: oodsS, queriesValues, friSamples, friDecommitments, traceDecommitments."""
        from . import proof_format
        return proof_format.size_breakdown_estimate(self.commitmentSchemeProof)
    size_breakdown_estimate = sizeBreakdownEstimate

    def toBytes(self) -> bytes:
        """The proof as bytes (format: tstwo_b200/proof_format.py; the reference defines none)."""
        from . import proof_format
        return proof_format.to_bytes(self.commitmentSchemeProof)
    to_bytes = toBytes

    @staticmethod
    def fromBytes(data: bytes) -> "StarkProof":
        from . import proof_format
        return StarkProof(proof_format.from_bytes(data))
    from_bytes = fromBytes

    def extractCompositionOodsEval(self) -> List[int]:
        """The last tree holds the 4 coordinate polynomials of the composition polynomial, each sampled at one point."""
        cols = self.sampledValues[-1]
        if len(cols) != SECURE_EXTENSION_DEGREE or any(len(c) != 1 for c in cols):
            raise ValueError("Unexpected sampled_values structure")
        return from_partial_evals([c[0] for c in cols])


def prove(components: Sequence, channel, commitment_scheme: CommitmentSchemeProver) -> StarkProof:
    """prover/index.ts:582-666."""
    if not isinstance(components, (list, tuple)):
        raise TypeError("prove: components must be an array")
    if len(components) == 0:
        raise ValueError("prove: components array cannot be empty")
    if channel is None:
        raise TypeError("prove: channel is required")
    if commitment_scheme is None:
        raise TypeError("prove: commitmentScheme is required")
    if len(commitment_scheme.trees) <= PREPROCESSED_TRACE_IDX:
        raise ValueError("prove: preprocessed trace not found in commitment scheme")
    n_preprocessed = len(commitment_scheme.trees[PREPROCESSED_TRACE_IDX].polynomials)
    provers = ComponentProvers(components, n_preprocessed)
    trace = Trace(commitment_scheme.polynomials(), commitment_scheme.evaluations())

    # evaluate and commit on the composition polynomial
    random_coeff = channel.draw_felt()
    composition = provers.computeCompositionPolynomial(random_coeff, trace, commitment_scheme.twiddles)
    builder = commitment_scheme.tree_builder()
    builder.extend_polys(composition.into_coordinate_polys())
    builder.commit(channel)

    oods_point = get_random_point(channel)
    sample_points = provers.maskPoints(oods_point)
    while len(sample_points) < len(commitment_scheme.trees) - 1:
        sample_points.append([])
    sample_points.append([[oods_point] for _ in range(SECURE_EXTENSION_DEGREE)])

    proof = StarkProof.create(commitment_scheme.prove_values(sample_points, channel))

    # sanity check: the committed composition polynomial at the point equals the constraints at the sampled mask values
    if proof.extractCompositionOodsEval() != provers.evalCompositionPolynomialAtPoint(oods_point, proof.sampledValues, random_coeff):
        raise ConstraintsNotSatisfied()
    return proof

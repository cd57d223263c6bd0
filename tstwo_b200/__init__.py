"""tstwo_b200: B200-native (sm_100a) CudaBackend for TStwo's LDE + Blake2s-commit hot path.

Layout: csrc/ (CUDA kernels + C ABI -> libtstwo_cuda.so), _lib.py (ctypes binding of include/tstwo_cuda.h),
circle.py (host index arithmetic), air.py (accumulators, wide-Fibonacci component), backend.py (mirror of the reference's backend interfaces), channel.py (Fiat-Shamir
channel), fri.py (FRI commit / decommit over the device folds and Merkle trees), pcs.py (prove_values), prover.py (prove)."""
from .backend import (CommitmentSchemeProver, CommitmentTreeProver, FriConfig, MerkleDecommitment, PcsConfig, TreeBuilder, TreeSubspan,
                      CudaBackend, CudaCircleEvaluation, CudaCirclePoly, CudaColumn, CudaFriOps,
                      CudaMerkleOps, LineEvaluation, MerkleProver, SecureColumnByCoords, SecureEvaluation, TwiddleTree,
                      accumulate, accumulate_quotients, fold_circle_into_line, fold_line)
from .air import (ColumnAccumulator, DomainEvaluationAccumulator, PointEvaluationAccumulator, SecureCirclePoly, WideFibonacciComponent,
                  WideFibonacciEval, generate_secure_powers, generate_wide_fibonacci_trace, widefib_constraints)
from .channel import Blake2sChannel, Blake2sMerkleChannel, ChannelTime
from .circle import CanonicCoset, CircleDomain, CirclePointIndex, Coset, LineDomain
from .fri import FriLayerProof, FriProof, FriProver, LinePoly, Queries
from .prover import ComponentProvers, ConstraintsNotSatisfied, StarkProof, Trace, prove
from .pcs import ColumnSampleBatch, CommitmentSchemeProof, PointSample, compute_fri_quotients, grind
from ._lib import TstwoError, TwiddleMismatch

__all__ = [n for n in dir() if not n.startswith("_")]

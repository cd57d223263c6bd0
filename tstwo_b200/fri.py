"""FRI commitment and decommitment over the device folds and Merkle kernels: the caller of fold_circle_into_line /
fold_line that the reference only sketches (packages/core/src/fri.ts:485-542, 618-785 carry placeholders for the Merkle
trees, the channel and the folding alphas).  This follows the control flow of those TS classes and fills the placeholders
in as Stwo's `FriProver` does (the Rust original the TS file quotes): roots and the last-layer polynomial go into the
Fiat-Shamir channel, every folding alpha is drawn from it.

Device work: folds (tstwo_fold_*), one Merkle tree per layer over the 4 coordinate columns (tstwo_merkle_commit), witness
gathers (tstwo_gather_words).  Host work, as in the reference: the channel, query bookkeeping, and the interpolation of
the last layer (a handful of values: `evaluation.toCpu().interpolate()`, fri.ts:729-731)."""
from __future__ import annotations

from typing import Dict, Iterable, List, Optional, Sequence

import numpy as np

from ._lib import check, dptr_array
from . import _lib
from .backend import (_L, CudaBackend, FriConfig, HashRows, LineEvaluation, MerkleDecommitment, MerkleProver, SecureColumnByCoords, SecureEvaluation,
                      TwiddleTree, fold_circle_into_line, fold_line)
from .channel import Blake2sChannel, Blake2sMerkleChannel
from .circle import Coset, LineDomain

P = 2147483647
CIRCLE_TO_LINE_FOLD_STEP = 1     # fri.ts:24
FOLD_STEP = 1                    # fri.ts:21
UPPER_BOUND_QUERY_BYTES = 4      # queries.ts:12


# ---------------------------------------------------------------------------------------------------- queries.ts
class Queries:
    """Sorted, de-duplicated query positions on a domain of 2^log_domain_size (queries.ts:14-172)."""

    def __init__(self, positions: Sequence[int], log_domain_size: int):
        if not isinstance(log_domain_size, int) or log_domain_size < 0:
            raise TypeError("logDomainSize must be a non-negative integer")
        if log_domain_size > 31:
            raise TypeError("logDomainSize must be at most 31 for JavaScript safety")
        prev = -1
        for i, p in enumerate(positions):
            if not isinstance(p, (int, np.integer)) or p < 0 or p >> log_domain_size:
                raise TypeError(f"Invalid position at index {i}: {p}")
            if p < prev:
                raise TypeError("Positions must be sorted in ascending order")
            prev = p
        self.positions = [int(p) for p in positions]
        self.log_domain_size = log_domain_size

    @staticmethod
    def generate(channel, log_domain_size: int, n_queries: int) -> "Queries":
        """queries.ts:50-82: little-endian u32 words of draw_random_bytes(), masked to the domain, until n distinct."""
        if not isinstance(n_queries, int) or n_queries < 0:
            raise TypeError("nQueries must be a non-negative integer")
        if not isinstance(log_domain_size, int) or log_domain_size < 0:
            raise TypeError("logDomainSize must be a non-negative integer")
        if log_domain_size < 31 and n_queries > 1 << log_domain_size:
            raise ValueError("more queries than domain points")
        found = set()
        mask = (1 << log_domain_size) - 1
        while len(found) < n_queries:
            b = channel.draw_random_bytes()
            for i in range(0, len(b) - UPPER_BOUND_QUERY_BYTES + 1, UPPER_BOUND_QUERY_BYTES):
                found.add(int.from_bytes(b[i:i + 4], "little") & mask)
                if len(found) == n_queries:
                    break
        return Queries(sorted(found), log_domain_size)

    @staticmethod
    def from_positions(positions: Sequence[int], log_domain_size: int) -> "Queries":
        return Queries(positions, log_domain_size)
    fromPositions = from_positions

    def fold(self, n_folds: int) -> "Queries":
        if not isinstance(n_folds, int) or n_folds < 0:
            raise TypeError("nFolds must be a non-negative integer")
        if n_folds > self.log_domain_size:
            raise ValueError("nFolds too large")
        return Queries(sorted({q >> n_folds for q in self.positions}), self.log_domain_size - n_folds)

    def __len__(self): return len(self.positions)
    def __iter__(self): return iter(self.positions)
    def __eq__(self, o): return self.positions == o.positions and self.log_domain_size == o.log_domain_size


def get_query_positions_by_log_size(queries: Queries, column_log_sizes: Iterable[int]) -> Dict[int, List[int]]:
    """fri.ts:470-481."""
    return {l: queries.fold(queries.log_domain_size - l).positions for l in sorted(set(column_log_sizes), reverse=True)}


def decommitment_positions(query_positions: Sequence[int], fold_step: int):
    """The position half of computeDecommitmentPositionsAndWitnessEvals (fri.ts:346-384): every position of every folding
    coset that holds a query, and which of them the verifier cannot compute itself (-> witness)."""
    positions, witness = [], []
    i = 0
    while i < len(query_positions):
        coset = query_positions[i] >> fold_step
        in_coset = set()
        while i < len(query_positions) and query_positions[i] >> fold_step == coset:
            in_coset.add(query_positions[i]); i += 1
        for pos in range(coset << fold_step, (coset + 1) << fold_step):
            positions.append(pos)
            if pos not in in_coset:
                witness.append(pos)
    return positions, witness


def compute_decommitment_positions_and_witness_evals(column: SecureColumnByCoords, query_positions: Sequence[int], fold_step: int):
    """fri.ts:346-384 on a device column: the witness evaluations are fetched by one gather launch."""
    positions, witness = decommitment_positions(query_positions, fold_step)
    n = column.len()
    for p in witness:
        if p >= n:
            raise IndexError("query position outside the column")
    evals = np.zeros((len(witness), 4), dtype=np.uint32)
    if witness:
        addrs = [c.ptr + 4 * p for p in witness for c in column.columns]
        check(_L().tstwo_gather_words(dptr_array(addrs), len(addrs), 1, evals.ctypes.data_as(_lib.C.POINTER(_lib.C.c_uint32))))
    return positions, [[int(x) for x in e] for e in evals]


# ---------------------------------------------------------------------------------------------------- poly/line.ts
def _bit_reverse(items: list) -> list:
    n = len(items)
    log = n.bit_length() - 1
    return [items[int(format(i, f"0{log}b")[::-1], 2) if log else 0] for i in range(n)]


class LinePoly:
    """poly/line.ts:127-238: coefficients (QM31 as 4 M31 words) in line_ifft's basis, stored bit-reversed."""

    def __init__(self, coeffs: Sequence[Sequence[int]]):
        n = len(coeffs)
        if n == 0 or n & (n - 1):
            raise ValueError("coeffs length must be power of two")
        self.coeffs = [[int(w) for w in c] for c in coeffs]

    @staticmethod
    def new(coeffs): return LinePoly(coeffs)

    @staticmethod
    def from_ordered_coefficients(coeffs): return LinePoly(_bit_reverse(list(coeffs)))
    fromOrderedCoefficients = from_ordered_coefficients

    def into_ordered_coefficients(self): return _bit_reverse(self.coeffs)
    intoOrderedCoefficients = into_ordered_coefficients
    def len(self): return len(self.coeffs)
    def __len__(self): return len(self.coeffs)
    def __eq__(self, o): return self.coeffs == o.coeffs


def line_interpolate(domain: LineDomain, coords: Sequence[np.ndarray]) -> LinePoly:
    """LineEvaluation.interpolate (poly/line.ts:311-328): bit-reverse, line_ifft (:354-390; the twiddles are M31, so the
    butterflies act on the 4 coordinates independently), scale by 1/n.  Host code, as in the reference."""
    n = domain.size()
    vals = [_bit_reverse([int(v) for v in c]) for c in coords]
    dom = domain
    while dom.size() > 1:
        size, half = dom.size(), dom.size() // 2
        inv = [pow(dom.at(i), P - 2, P) for i in range(half)]
        for v in vals:
            for start in range(0, n, size):
                for i in range(half):
                    a, b = v[start + i], v[start + i + half]
                    v[start + i], v[start + i + half] = (a + b) % P, (a - b) * inv[i] % P
        dom = dom.double()
    n_inv = pow(n, P - 2, P)
    return LinePoly([[vals[k][i] * n_inv % P for k in range(4)] for i in range(n)])


# ---------------------------------------------------------------------------------------------------- fri.ts provers
class FriLayerProof:
    """fri.ts:254-258."""

    def __init__(self, fri_witness, decommitment: MerkleDecommitment, commitment: bytes):
        self.fri_witness, self.decommitment, self.commitment = fri_witness, decommitment, commitment


class FriProof:
    """fri.ts:260-264."""

    def __init__(self, first_layer: FriLayerProof, inner_layers: List[FriLayerProof], last_layer_poly: LinePoly):
        self.first_layer, self.inner_layers, self.last_layer_poly = first_layer, inner_layers, last_layer_poly


def extract_coordinate_columns(columns: Sequence[SecureEvaluation]):
    """fri.ts:607-616."""
    return [c for col in columns for c in col.values.columns]


class FriFirstLayerProver:
    """fri.ts:485-517: all circle columns under one Merkle tree."""

    def __init__(self, columns: Sequence[SecureEvaluation], merkle_tree: MerkleProver):
        self.columns, self.merkle_tree = list(columns), merkle_tree

    @staticmethod
    def new(columns: Sequence[SecureEvaluation]) -> "FriFirstLayerProver":
        return FriFirstLayerProver(columns, MerkleProver.commit(CudaBackend.merkleOps, extract_coordinate_columns(columns)))

    def column_log_sizes(self): return {c.domain.logSize() for c in self.columns}
    columnLogSizes = column_log_sizes
    def max_column_log_size(self): return max(self.column_log_sizes())
    maxColumnLogSize = max_column_log_size

    def decommit(self, queries: Queries) -> FriLayerProof:
        if queries.log_domain_size != self.max_column_log_size():
            raise ValueError("queries are not on the largest column's domain")
        witness, by_log = [], {}
        for col in self.columns:
            log = col.domain.logSize()
            col_queries = queries.fold(queries.log_domain_size - log)
            pos, w = compute_decommitment_positions_and_witness_evals(col.values, col_queries.positions, CIRCLE_TO_LINE_FOLD_STEP)
            by_log[log] = pos
            witness += w
        _, dec = self.merkle_tree.decommit(by_log, extract_coordinate_columns(self.columns))
        return FriLayerProof(witness, dec, self.merkle_tree.root())


class FriInnerLayerProver:
    """fri.ts:519-542: one line evaluation under its own Merkle tree (4 coordinate columns)."""

    def __init__(self, evaluation: LineEvaluation, merkle_tree: MerkleProver):
        self.evaluation, self.merkle_tree = evaluation, merkle_tree

    @staticmethod
    def new(evaluation: LineEvaluation) -> "FriInnerLayerProver":
        return FriInnerLayerProver(evaluation, MerkleProver.commit(CudaBackend.merkleOps, evaluation.values.columns))

    def decommit(self, queries: Queries) -> FriLayerProof:
        pos, w = compute_decommitment_positions_and_witness_evals(self.evaluation.values, queries.positions, FOLD_STEP)
        _, dec = self.merkle_tree.decommit({self.evaluation.domain().logSize(): pos}, self.evaluation.values.columns)
        return FriLayerProof(w, dec, self.merkle_tree.root())


class FriProver:
    """fri.ts:618-785."""

    def __init__(self, config: FriConfig, first_layer: FriFirstLayerProver, inner_layers: List[FriInnerLayerProver], last_layer_poly: LinePoly):
        self.config, self.first_layer, self.inner_layers, self.last_layer_poly = config, first_layer, inner_layers, last_layer_poly

    @staticmethod
    def commit(channel, config: FriConfig, columns: Sequence[SecureEvaluation], twiddles: Optional[TwiddleTree] = None) -> "FriProver":
        """`columns` in descending order by size, at most one per size, on canonic domains (fri.ts:645-675)."""
        if len(columns) == 0:
            raise ValueError("no columns")
        if not all(c.domain.isCanonic() for c in columns):
            raise ValueError("not canonic")
        if any(a.len() <= b.len() for a, b in zip(columns, columns[1:])):
            raise ValueError("column sizes not decreasing")
        first_layer = FriFirstLayerProver.new(columns)
        Blake2sMerkleChannel.mix_root(channel, first_layer.merkle_tree.root())
        inner_layers, last_eval = FriProver._commit_inner_layers(channel, config, columns, twiddles)
        last_layer_poly = FriProver._commit_last_layer(channel, config, last_eval)
        return FriProver(config, first_layer, inner_layers, last_layer_poly)

    @staticmethod
    def _commit_inner_layers(channel, config, columns, twiddles):
        """fri.ts:676-716 with the channel in place of the placeholders; a circle column joins the layer whose size its
        folded size equals and is only consumed then (Stwo's `next_if`; the TS loop drops a non-matching column).
        The circle column is folded with the alpha of the layer it joins, as the reference's text does (fri.ts:704-711,
        a port of Stwo 0.1.1's commit_inner_layers).

        The whole loop is ONE library call (tstwo_fri_commit_layers): trees, mix_root, draw_felt and folds run back to back
        on the device with the transcript in device memory; the channel object is updated from the state that comes back.
        FriProver.per_layer_calls = True takes the layer-by-layer form (one host round trip per layer) instead."""
        folded_size = lambda v: v.len() >> CIRCLE_TO_LINE_FOLD_STEP
        first_log = folded_size(columns[0]).bit_length() - 1
        layer_evaluation = LineEvaluation.newZero(LineDomain.new(Coset.half_odds(first_log)))
        pending = list(columns)
        layers: List[FriInnerLayerProver] = []
        alpha = channel.draw_felt()
        fold_circle_into_line(layer_evaluation, pending.pop(0), alpha, twiddles)
        if not FriProver.per_layer_calls and isinstance(channel, Blake2sChannel):
            return FriProver._commit_inner_layers_device(channel, config, layer_evaluation, pending, twiddles)
        while layer_evaluation.len() > config.last_layer_domain_size():
            layer = FriInnerLayerProver.new(layer_evaluation)
            Blake2sMerkleChannel.mix_root(channel, layer.merkle_tree.root())
            alpha = channel.draw_felt()
            layer_evaluation = fold_line(layer.evaluation, alpha, twiddles)
            if pending and folded_size(pending[0]) == layer_evaluation.len():
                fold_circle_into_line(layer_evaluation, pending.pop(0), alpha, twiddles)
            layers.append(layer)
        if pending:
            raise ValueError("columns left over after the last layer")
        return layers, layer_evaluation

    per_layer_calls = False

    @staticmethod
    def _commit_inner_layers_device(channel, config, layer_evaluation, pending, twiddles):
        import ctypes as C
        from . import _lib
        from .backend import CudaColumn, SecureColumnByCoords, MerkleProver
        L = _lib.load()
        first_log = layer_evaluation.len().bit_length() - 1
        last_log = config.last_layer_domain_size().bit_length() - 1
        n_layers = max(0, first_log - last_log)
        folded_log = lambda v: v.domain.logSize() - CIRCLE_TO_LINE_FOLD_STEP
        if any(folded_log(c) >= first_log or folded_log(c) < last_log for c in pending):
            raise ValueError("columns left over after the last layer")
        if n_layers == 0:
            if pending:
                raise ValueError("columns left over after the last layer")
            return [], layer_evaluation
        n_c = len(pending)
        cptrs = _lib.dptr_array([col.ptr for c in pending for col in c.values.columns])
        clogs = (C.c_int * max(1, n_c))(*[c.domain.logSize() for c in pending])
        cinits = _lib.u32_array([c.domain.halfCoset.initial_index.value for c in pending])
        st = channel._digest + channel._n_challenges.to_bytes(4, "little") + channel._n_sent.to_bytes(4, "little")
        q = list(channel._base_queue)
        state = (C.c_uint32 * 19)(*(list(int.from_bytes(st[4 * i:4 * i + 4], "little") for i in range(10)) + [len(q)] + q + [0] * (8 - len(q))))
        evals = (C.c_uint64 * (4 * n_layers))()
        trees = (C.c_uint64 * n_layers)()
        roots = (C.c_uint8 * (32 * n_layers))()
        alphas = (C.c_uint32 * (4 * n_layers))()
        n_out = C.c_int()
        dom = layer_evaluation.domain()
        _lib.check(L.tstwo_fri_commit_layers(layer_evaluation.values.ptrs(), first_log, dom.coset().initial_index.value, last_log,
                                             cptrs, clogs, cinits, n_c, twiddles.handle if twiddles is not None else 0, state,
                                             evals, trees, roots, alphas, C.byref(n_out)))
        assert n_out.value == n_layers
        # the transcript continues from the device's state
        channel._digest = b"".join(int(state[i]).to_bytes(4, "little") for i in range(8))
        channel._n_challenges, channel._n_sent = int(state[8]), int(state[9])
        channel._base_queue = [int(state[11 + i]) for i in range(int(state[10]))]
        layers: List[FriInnerLayerProver] = []
        cur = layer_evaluation
        coset = dom.coset()
        for l in range(n_layers):
            log = first_log - l
            buf = CudaColumn(int(trees[l]), 8 * ((2 << log) - 1))
            tree_layers = [buf.slice(8 * ((1 << k) - 1), 8 << k) for k in range(log + 1)]
            layers.append(FriInnerLayerProver(cur, MerkleProver(tree_layers, bytes(roots[32 * l:32 * l + 32]))))
            coset = coset.double()
            stride = max(4, 1 << (log - 1))
            block = CudaColumn(int(evals[4 * l]), 4 * stride)              # the four coordinates are one block (owner)
            cols = [block.slice(k * stride, 1 << (log - 1)) for k in range(4)]
            cur = LineEvaluation.new(LineDomain.new(coset), SecureColumnByCoords(cols))
        FriProver.last_alphas = [[int(alphas[4 * l + k]) for k in range(4)] for l in range(n_layers)]
        return layers, cur

    @staticmethod
    def _commit_last_layer(channel, config, evaluation: LineEvaluation) -> LinePoly:
        """fri.ts:718-754."""
        if evaluation.len() != config.last_layer_domain_size():
            raise ValueError("last layer domain size mismatch")
        coeffs = line_interpolate(evaluation.domain(), evaluation.values.to_cpu()).into_ordered_coefficients()
        bound = 1 << config.log_last_layer_degree_bound
        if any(any(c) for c in coeffs[bound:]):
            raise ValueError("invalid degree")
        poly = LinePoly.from_ordered_coefficients(coeffs[:bound])
        channel.mix_felts(poly.coeffs)
        return poly

    def decommit(self, channel):
        """fri.ts:759-767: (proof, query positions per column log size)."""
        queries = Queries.generate(channel, self.first_layer.max_column_log_size(), self.config.n_queries)
        by_log = get_query_positions_by_log_size(queries, self.first_layer.column_log_sizes())
        return self.decommit_on_queries(queries), by_log

    def decommit_on_queries(self, queries: Queries) -> FriProof:
        """fri.ts:769-785."""
        first = self.first_layer.decommit(queries)
        layer_queries = queries.fold(CIRCLE_TO_LINE_FOLD_STEP)
        if FriProver.per_layer_calls or not self.inner_layers or not all(self._tree_is_one_buffer(l) for l in self.inner_layers):
            inner = []
            for layer in self.inner_layers:
                inner.append(layer.decommit(layer_queries))
                layer_queries = layer_queries.fold(FOLD_STEP)
            return FriProof(first, inner, self.last_layer_poly)
        return FriProof(first, self._decommit_inner_layers_device(layer_queries), self.last_layer_poly)

    @staticmethod
    def _tree_is_one_buffer(layer: "FriInnerLayerProver") -> bool:
        ls = layer.merkle_tree.layers
        return all(ls[k].ptr == ls[0].ptr + 32 * ((1 << k) - 1) for k in range(len(ls)))

    def _decommit_inner_layers_device(self, layer_queries: Queries) -> List[FriLayerProof]:
        """All inner layers in ONE library call (tstwo_fri_decommit_layers): positions, FRI witnesses and Merkle walks of every
        layer on the host inside the library, two gather launches and one synchronisation for all of them."""
        import ctypes as C
        L = _L()
        n = len(self.inner_layers)
        logs = [l.evaluation.domain().logSize() for l in self.inner_layers]
        trees = dptr_array([l.merkle_tree.layers[0].ptr for l in self.inner_layers])
        evals = dptr_array([c.ptr for l in self.inner_layers for c in l.evaluation.values.columns])
        q = list(layer_queries.positions)
        nq = len(q)
        cap_w, cap_h = max(1, nq * n), max(1, 2 * nq * sum(logs))
        wit = np.zeros((cap_w, 4), dtype=np.uint32)
        hashes = np.zeros((cap_h, 32), dtype=np.uint8)
        colw = np.zeros(max(1, 8 * nq * n), dtype=np.uint32)
        counts = (C.c_uint64 * (3 * n))()
        check(L.tstwo_fri_decommit_layers(n, trees, evals, (C.c_int * n)(*logs), _lib.u32_array(q if q else [0]), nq,
                                          wit.ctypes.data_as(_lib.u32p), cap_w, hashes.ctypes.data_as(C.POINTER(C.c_uint8)), cap_h,
                                          colw.ctypes.data_as(_lib.u32p), colw.size, counts))
        out, wi, hi, ci = [], 0, 0, 0
        for l, layer in enumerate(self.inner_layers):
            nw, nh, nc = int(counts[3 * l]), int(counts[3 * l + 1]), int(counts[3 * l + 2])
            dec = MerkleDecommitment(HashRows(hashes[hi:hi + nh]), colw[ci:ci + nc].copy())
            out.append(FriLayerProof(wit[wi:wi + nw].tolist(), dec, layer.merkle_tree.root()))
            wi, hi, ci = wi + nw, hi + nh, ci + nc
        return out
    decommitOnQueries = decommit_on_queries

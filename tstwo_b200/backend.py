"""CudaBackend: host-side mirror of the reference's backend interfaces for the LDE + Blake2s-commit hot path,
over the C ABI of libtstwo_cuda.so (include/tstwo_cuda.h).

Mirrors (names, argument meaning, error behaviour), paths relative to packages/core/src of teddyjfpender/tstwo:
  Column / Backend            backend/index.ts:20-74, backend/cpu/index.ts:18-157
  CpuCircleEvaluation / Poly  backend/cpu/circle.ts:17-208, poly/circle/evaluation.ts:98-177, poly/circle/poly.ts:9-73
  PolyOps batch entry points  poly/circle/ops.ts:20-101
  TwiddleTree                 poly/twiddles.ts:14-29
  MerkleOps / MerkleProver    vcs/ops.ts:16-26, vcs/prover.ts:6-30,111-113
  FriOps                      fri.ts:93-110, backend/cpu/fri.ts:23-226
  accumulate / quotients      backend/cpu/accumulation.ts:38-63, backend/cpu/quotients.ts:52-178
  CommitmentTreeProver        pcs/prover.ts:215-235 (Rust text; never ported to TS)

The reference's host language (TypeScript on Bun) is not available in the build image; the same classes are
provided for Bun in ts/ (see INTEGRATION.md).  There is NO CPU fallback here: every method ends in a C-ABI call
that fails loudly without a B200.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence

import numpy as np

from . import _lib
from ._lib import FLAG_QUIRKS, TstwoError, TwiddleMismatch, check, dptr_array, u32_array
from .channel import Blake2sChannel, Blake2sMerkleChannel
from .circle import CanonicCoset, CircleDomain, CirclePointIndex, Coset, LineDomain

P = 2147483647


def _L():
    return _lib.init(CudaBackend.device)


class CudaColumn:
    """Column<BaseField> held in device memory (backend/index.ts:53-74).  Values are canonical u32 in [0, P)."""

    def __init__(self, ptr: int, length: int, owner=None):
        self.ptr = ptr
        self._len = length
        self._owner = owner       # keeps a parent allocation (or a torch tensor) alive
        self._owns = owner is None

    # -- construction ---------------------------------------------------------------------------------------
    @staticmethod
    def uninitialized(length: int) -> "CudaColumn":
        p = C.c_uint64()
        check(_L().tstwo_malloc(4 * length, C.byref(p)))
        return CudaColumn(p.value, length)

    @staticmethod
    def zeros(length: int) -> "CudaColumn":
        col = CudaColumn.uninitialized(length)
        check(_L().tstwo_memset_zero(col.ptr, 4 * length))
        return col

    @staticmethod
    def from_cpu(values) -> "CudaColumn":
        a = np.ascontiguousarray(np.asarray(values, dtype=np.uint32))
        if a.size and int(a.max()) >= P:
            raise ValueError("from_u32_unchecked: val must be an integer in [0, P)")   # fields/m31.ts:81-86
        col = CudaColumn.uninitialized(a.size)
        if a.size:
            check(_L().tstwo_upload(col.ptr, a.ctypes.data, 4 * a.size))
        return col

    @staticmethod
    def view(ptr: int, length: int, owner) -> "CudaColumn":
        return CudaColumn(ptr, length, owner)

    # -- Column<T> ------------------------------------------------------------------------------------------
    def len(self) -> int: return self._len
    def __len__(self) -> int: return self._len
    def isEmpty(self) -> bool: return self._len == 0

    def toCpu(self) -> np.ndarray:
        out = np.empty(self._len, dtype=np.uint32)
        if self._len:
            check(_L().tstwo_download(out.ctypes.data, self.ptr, 4 * self._len))
        return out
    to_cpu = toCpu

    def at(self, index: int) -> int:
        if index < 0 or index >= self._len:
            raise IndexError(f"Index {index} out of bounds for length {self._len}")
        v = C.c_uint32()
        check(_L().tstwo_download(C.byref(v), self.ptr + 4 * index, 4))
        return v.value

    def set(self, index: int, value: int) -> None:
        if index < 0 or index >= self._len:
            raise IndexError(f"Index {index} out of bounds for length {self._len}")
        v = C.c_uint32(int(value))
        check(_L().tstwo_upload(self.ptr + 4 * index, C.byref(v), 4))

    def clone(self) -> "CudaColumn":
        col = CudaColumn.uninitialized(self._len)
        check(_L().tstwo_copy(col.ptr, self.ptr, 4 * self._len))
        return col

    def slice(self, off: int, length: int) -> "CudaColumn":
        return CudaColumn(self.ptr + 4 * off, length, owner=self)

    def __del__(self):
        if getattr(self, "_owns", False) and self.ptr and _lib is not None and _lib._lib is not None:   # module globals go first at exit
            try:
                _lib._lib.tstwo_free(self.ptr)
            except Exception:
                pass
            self.ptr = 0


class SecureColumnByCoords:
    """fields/secure_columns.ts:124-217: a SecureField column as 4 base-field coordinate columns."""

    def __init__(self, columns: Sequence[CudaColumn]):
        assert len(columns) == 4
        self.columns = list(columns)

    # The four coordinates are the quarters of ONE block: evenly spaced columns are what the bulk-copy (TMA) kernels take --
    # MerkleProver.commit of a secure column then runs the leaf layer at 25.7 instead of 19 G compressions/s.
    @staticmethod
    def _of_block(block: CudaColumn, length: int):
        return SecureColumnByCoords([block.slice(k * length, length) for k in range(4)])

    @staticmethod
    def zeros(length: int):
        if length % 4:                                            # tiny columns: keep every coordinate 16-byte aligned
            return SecureColumnByCoords([CudaColumn.zeros(length) for _ in range(4)])
        return SecureColumnByCoords._of_block(CudaColumn.zeros(4 * length), length)

    @staticmethod
    def uninitialized(length: int):
        if length % 4:
            return SecureColumnByCoords([CudaColumn.uninitialized(length) for _ in range(4)])
        return SecureColumnByCoords._of_block(CudaColumn.uninitialized(4 * length), length)

    @staticmethod
    def from_cpu(coords):
        a = np.stack([np.asarray(c, dtype=np.uint32) for c in coords])
        if a.shape[0] != 4:
            raise ValueError("a secure column has 4 coordinate columns")
        if a.shape[1] % 4:
            return SecureColumnByCoords([CudaColumn.from_cpu(c) for c in a])
        return SecureColumnByCoords._of_block(CudaColumn.from_cpu(a.reshape(-1)), a.shape[1])

    def len(self): return self.columns[0].len()
    def at(self, i): return [c.at(i) for c in self.columns]
    def set(self, i, v):
        for c, x in zip(self.columns, v):
            c.set(i, x)
    def to_cpu(self): return [c.toCpu() for c in self.columns]
    def ptrs(self): return dptr_array([c.ptr for c in self.columns])


class TwiddleTree:
    """poly/twiddles.ts:14-29; buffers live on the device in the reference's layout."""

    def __init__(self, root_coset: Coset, handle: int):
        self.rootCoset = root_coset
        self.handle = handle

    def _info(self):
        tw, itw = C.c_uint64(), C.c_uint64()
        check(_L().tstwo_twiddles_info(self.handle, None, None, C.byref(tw), C.byref(itw)))
        return tw.value, itw.value

    @property
    def twiddles(self) -> CudaColumn:
        return CudaColumn.view(self._info()[0], self.rootCoset.size(), self)

    @property
    def itwiddles(self) -> CudaColumn:
        return CudaColumn.view(self._info()[1], self.rootCoset.size(), self)

    def __del__(self):
        if self.handle and _lib is not None and _lib._lib is not None:
            try:
                _lib._lib.tstwo_twiddles_free(self.handle)
            except Exception:
                pass
            self.handle = 0


def _flags():
    return FLAG_QUIRKS if CudaBackend.quirks else 0


class CudaCirclePoly:
    """CpuCirclePoly (backend/cpu/circle.ts:39-208) on the device.  coeffs in the bit-reversed tensor basis."""

    def __init__(self, coeffs: CudaColumn):
        n = coeffs.len()
        if n == 0 or n & (n - 1):
            raise ValueError("coeffs length must be power of two")
        self.coeffs = coeffs

    @staticmethod
    def new(coeffs) -> "CudaCirclePoly":
        return CudaCirclePoly(coeffs if isinstance(coeffs, CudaColumn) else CudaColumn.from_cpu(coeffs))

    def logSize(self) -> int: return self.coeffs.len().bit_length() - 1
    log_size = logSize

    @staticmethod
    def precomputeTwiddles(coset: Coset) -> TwiddleTree:
        return CudaBackend.precomputeTwiddles(coset)

    @staticmethod
    def eval_at_point(poly: "CudaCirclePoly", point) -> List[int]:
        """point = (x[4], y[4]) CirclePoint<SecureField>; returns QM31 as [a,b,c,d] (circle.ts:52-69)."""
        out = (C.c_uint32 * 4)()
        check(_L().tstwo_eval_at_point(poly.coeffs.ptr, poly.logSize(), u32_array(point[0]), u32_array(point[1]), out))
        return list(out)

    @staticmethod
    def eval_many_at_point(polys: Sequence["CudaCirclePoly"], point) -> List[List[int]]:
        """eval_at_point of several polynomials at one point: one batch of launches per polynomial size (tstwo_eval_at_point_many)."""
        out: List[Optional[List[int]]] = [None] * len(polys)
        by_log = {}
        for i, p in enumerate(polys):
            by_log.setdefault(p.logSize(), []).append(i)
        for log, idxs in by_log.items():
            res = (C.c_uint32 * (4 * len(idxs)))()
            check(_L().tstwo_eval_at_point_many(dptr_array([polys[i].coeffs.ptr for i in idxs]), len(idxs), log, u32_array(point[0]),
                                                u32_array(point[1]), res))
            for k, i in enumerate(idxs):
                out[i] = list(res[4 * k:4 * k + 4])
        return out

    @staticmethod
    def extend(poly: "CudaCirclePoly", log_size: int) -> "CudaCirclePoly":
        if log_size < poly.logSize():
            raise ValueError("log size too small")
        col = CudaColumn.zeros(1 << log_size)
        check(_L().tstwo_copy(col.ptr, poly.coeffs.ptr, 4 * poly.coeffs.len()))
        return CudaCirclePoly(col)

    @staticmethod
    def evaluate(poly: "CudaCirclePoly", domain: CircleDomain, twiddles: TwiddleTree) -> "CudaCircleEvaluation":
        return CudaBackend.evaluatePolynomialsOnDomain([poly], domain, twiddles)[0]

    # instance-level conveniences of CirclePoly (poly/circle/poly.ts:34-54)
    def evaluateWithTwiddles(self, domain, twiddles): return CudaCirclePoly.evaluate(self, domain, twiddles)
    def evaluate_(self, domain): return CudaCirclePoly.evaluate(self, domain, CudaBackend.precomputeTwiddles(domain.halfCoset))


class CudaCircleEvaluation:
    """CpuCircleEvaluation (backend/cpu/circle.ts:17-37; poly/circle/evaluation.ts:98-177), bit-reversed order."""

    def __init__(self, domain: CircleDomain, values: CudaColumn):
        if domain.size() != values.len():
            raise ValueError("domain size must match values length")   # evaluation.ts:106-108
        self.domain = domain
        self.values = values

    @staticmethod
    def new(domain: CircleDomain, values) -> "CudaCircleEvaluation":
        return CudaCircleEvaluation(domain, values if isinstance(values, CudaColumn) else CudaColumn.from_cpu(values))

    @staticmethod
    def bitReverseColumn(col: CudaColumn) -> None: CudaBackend.bitReverseColumn(col)

    @staticmethod
    def precomputeTwiddles(coset: Coset) -> TwiddleTree: return CudaBackend.precomputeTwiddles(coset)

    @staticmethod
    def to_cpu(values: CudaColumn) -> np.ndarray: return values.toCpu()

    def interpolate(self) -> CudaCirclePoly:
        return self.interpolateWithTwiddles(CudaBackend.precomputeTwiddles(self.domain.halfCoset))

    def interpolateWithTwiddles(self, twiddles: TwiddleTree) -> CudaCirclePoly:
        return CudaBackend.interpolateColumns([self], twiddles)[0]

    def bitReverse(self) -> "CudaCircleEvaluation":
        col = self.values.clone()
        CudaBackend.bitReverseColumn(col)
        return CudaCircleEvaluation(self.domain, col)


class CudaMerkleOps:
    """MerkleOps<Blake2sHash> (vcs/ops.ts:16-26) with Blake2sMerkleHasher.hashNode semantics."""

    def commitOnLayer(self, logSize: int, prevLayer: Optional[CudaColumn], columns: Sequence[CudaColumn]) -> CudaColumn:
        for c in columns:
            if c.len() != 1 << logSize:
                raise ValueError("column length does not match layer size")
        if prevLayer is not None and prevLayer.len() != 16 << logSize:
            raise ValueError("Child hashes must be 32 bytes long.")
        out = CudaColumn.uninitialized(8 << logSize)     # 32 bytes per hash
        check(_L().tstwo_merkle_commit_layer(logSize, prevLayer.ptr if prevLayer is not None else 0,
                                             dptr_array([c.ptr for c in columns]), len(columns), out.ptr))
        return out
    commit_on_layer = commitOnLayer


class MerkleProver:
    """vcs/prover.ts:6-30,111-113.  layers[0] is the root layer; each layer is a device buffer of 32-byte hashes."""

    def __init__(self, layers: List[CudaColumn], root: bytes):
        self.layers = layers
        self._root = root

    @staticmethod
    def commit(ops: Optional[CudaMerkleOps], columns: Sequence[CudaColumn]) -> "MerkleProver":
        cols = list(columns)
        logs = []
        for c in cols:
            n = c.len()
            if n == 0 or n & (n - 1):
                raise ValueError("length is not power of two")
            logs.append(n.bit_length() - 1)
        max_log = max(logs) if logs else 0
        buf = CudaColumn.uninitialized(8 * ((2 << max_log) - 1))
        root = (C.c_uint8 * 32)()
        check(_L().tstwo_merkle_commit(dptr_array([c.ptr for c in cols]), (C.c_int * max(1, len(logs)))(*logs), len(cols),
                                       buf.ptr, root))
        layers = [buf.slice(8 * ((1 << k) - 1), 8 << k) for k in range(max_log + 1)]
        return MerkleProver(layers, bytes(root))

    def root(self) -> bytes: return self._root

    def layer_bytes(self, k: int) -> np.ndarray:
        return self.layers[k].toCpu().view(np.uint8).reshape(-1, 32)

    def decommit(self, queriesPerLogSize, columns: Sequence[CudaColumn]):
        """vcs/prover.ts:32-109: (queried values, MerkleDecommitment) for sorted query positions per column log size.
        Which hashes and values go into the witness depends on the queries only, so the tree walk runs on the host and
        the witness itself is collected from device memory by two gather launches (tstwo_gather_words)."""
        max_log = len(self.layers) - 1
        base = self.layers[0].ptr
        if all(self.layers[k].ptr == base + 32 * ((1 << k) - 1) for k in range(max_log + 1)):
            return self._decommit_native(queriesPerLogSize, list(columns), base, max_log)
        cols = sorted(columns, key=lambda c: -c.len())          # stable, as Array.prototype.sort in the reference
        col_i = 0
        hash_addrs: List[int] = []                              # device address of every witness hash, in order
        value_addrs: List[int] = []                             # ... of every column value, in node order
        value_is_query: List[bool] = []
        last: List[int] = []
        for log in range(len(self.layers) - 1, -1, -1):
            layer_cols = []
            while col_i < len(cols) and cols[col_i].len() == 1 << log:
                layer_cols.append(cols[col_i]); col_i += 1
            child = self.layers[log + 1] if log + 1 < len(self.layers) else None
            direct = list(queriesPerLogSize.get(log, ()) or ())
            pi = di = 0
            nxt: List[int] = []
            while pi < len(last) or di < len(direct):
                cand = []                                        # nextDecommitmentNode, vcs/utils.ts:40-57
                if pi < len(last): cand.append(last[pi] >> 1)
                if di < len(direct): cand.append(direct[di])
                node = min(cand)
                if child is not None:
                    for ch in (2 * node, 2 * node + 1):
                        if pi < len(last) and last[pi] == ch:
                            pi += 1                              # the verifier recomputes this child itself
                        else:
                            if ch >= child.len() // 8:
                                raise IndexError(f"MerkleProver.decommit: child {ch} is outside layer {log + 1}")
                            hash_addrs.append(child.ptr + 32 * ch)
                queried = di < len(direct) and direct[di] == node
                if queried:
                    di += 1
                if node >= 1 << log:
                    raise IndexError(f"MerkleProver.decommit: query {node} is outside layer {log}")
                for c in layer_cols:
                    value_addrs.append(c.ptr + 4 * node)
                    value_is_query.append(queried)
                nxt.append(node)
            last = nxt
        hashes = np.zeros((len(hash_addrs), 8), dtype=np.uint32)
        values = np.zeros(len(value_addrs), dtype=np.uint32)
        if hash_addrs:
            check(_L().tstwo_gather_words(dptr_array(hash_addrs), len(hash_addrs), 8, hashes.ctypes.data_as(C.POINTER(C.c_uint32))))
        if value_addrs:
            check(_L().tstwo_gather_words(dptr_array(value_addrs), len(value_addrs), 1, values.ctypes.data_as(C.POINTER(C.c_uint32))))
        mask = np.array(value_is_query, dtype=bool)
        return values[mask], MerkleDecommitment([h.tobytes() for h in hashes], values[~mask])


    def _decommit_native(self, queriesPerLogSize, columns, base: int, max_log: int):
        """The same walk inside the library (tstwo_merkle_decommit): one call, two gather launches, one synchronisation."""
        offsets, flat = [0], []
        for log in range(max_log + 1):
            direct = list(queriesPerLogSize.get(log, ()) or ())
            if direct and max(direct) >> log:
                raise IndexError(f"MerkleProver.decommit: query {max(direct)} is outside layer {log}")
            flat.extend(direct); offsets.append(len(flat))
        logs = []
        for c in columns:
            n = c.len()
            if n == 0 or n & (n - 1) or n.bit_length() - 1 > max_log:
                raise ValueError("column does not fit the tree")
            logs.append(n.bit_length() - 1)
        L = _L()
        args = (base, max_log, dptr_array([c.ptr for c in columns]), (C.c_int * max(1, len(logs)))(*logs), len(columns),
                (C.c_int * len(offsets))(*offsets), u32_array(flat if flat else [0]))
        counts = (C.c_uint64 * 3)()
        check(L.tstwo_merkle_decommit(*args, None, None, None, counts))
        queried = np.zeros(max(1, counts[0]), dtype=np.uint32)
        hashes = np.zeros((max(1, counts[1]), 32), dtype=np.uint8)
        witness = np.zeros(max(1, counts[2]), dtype=np.uint32)
        check(L.tstwo_merkle_decommit(*args, queried.ctypes.data_as(C.POINTER(C.c_uint32)), hashes.ctypes.data_as(C.POINTER(C.c_uint8)),
                                      witness.ctypes.data_as(C.POINTER(C.c_uint32)), counts))
        return queried[:counts[0]], MerkleDecommitment([h.tobytes() for h in hashes[:counts[1]]], witness[:counts[2]])


class HashRows:
    """A read-only sequence of 32-byte hashes over the rows of an (n, 32) uint8 array: what a list of `bytes` would be, without
    materialising tens of thousands of bytes objects per proof (the inner FRI layers of a 70-query proof carry ~30 000 witness hashes)."""

    def __init__(self, rows: np.ndarray):
        self._rows = rows

    def __len__(self): return self._rows.shape[0]

    def __getitem__(self, i):
        if isinstance(i, slice):
            return HashRows(self._rows[i])
        return self._rows[i].tobytes()

    def __iter__(self): return (r.tobytes() for r in self._rows)

    def __eq__(self, other):
        if isinstance(other, HashRows):
            return self._rows.shape == other._rows.shape and bool((self._rows == other._rows).all())
        try:
            return len(other) == len(self) and all(bytes(a) == b for a, b in zip(other, self))
        except TypeError:
            return NotImplemented

    def __repr__(self): return f"HashRows({len(self)} hashes)"


class MerkleDecommitment:
    """vcs/verifier.ts:5-8."""

    def __init__(self, hashWitness: List[bytes], columnWitness):
        self.hashWitness = hashWitness
        self.columnWitness = columnWitness
    hash_witness = property(lambda self: self.hashWitness)
    column_witness = property(lambda self: self.columnWitness)


class LineEvaluation:
    """poly/line.ts:241-329 over device SecureColumnByCoords."""

    def __init__(self, domain: LineDomain, values: SecureColumnByCoords):
        if domain.size() != values.len():
            raise ValueError("domain size must match values length")
        self._domain = domain
        self.values = values

    @staticmethod
    def new(domain, values): return LineEvaluation(domain, values)

    @staticmethod
    def newZero(domain): return LineEvaluation(domain, SecureColumnByCoords.zeros(domain.size()))
    new_zero = newZero

    def len(self): return self._domain.size()
    def domain(self): return self._domain


class SecureEvaluation:
    """poly/circle/secure_poly.ts:46-81 (bit-reversed order) over device SecureColumnByCoords."""

    def __init__(self, domain: CircleDomain, values: SecureColumnByCoords):
        if domain.size() != values.len():
            raise ValueError("domain size must match values length")
        self.domain = domain
        self.values = values

    def len(self): return self.domain.size()


class CudaFriOps:
    """FriOps (fri.ts:93-110; backend/cpu/fri.ts:175-226)."""

    def foldLine(self, eval_: LineEvaluation, alpha, twiddles: Optional[TwiddleTree] = None) -> LineEvaluation:
        n = eval_.len()
        if n < 2:
            raise ValueError("fold_line: Evaluation too small, must have at least 2 elements.")
        dom = eval_.domain()
        out = SecureColumnByCoords.uninitialized(n // 2)
        check(_L().tstwo_fold_line(eval_.values.ptrs(), dom.logSize(), dom.coset().initial_index.value, u32_array(alpha),
                                   out.ptrs(), twiddles.handle if twiddles else 0))
        return LineEvaluation(dom.double(), out)
    fold_line = foldLine

    def foldCircleIntoLine(self, dst: LineEvaluation, src: SecureEvaluation, alpha, twiddles: Optional[TwiddleTree] = None) -> None:
        if (src.domain.size() >> 1) != dst.len():
            raise ValueError("fold_circle_into_line: Length mismatch between src and dst after considering fold step.")
        check(_L().tstwo_fold_circle_into_line(dst.values.ptrs(), src.values.ptrs(), src.domain.logSize(),
                                               src.domain.halfCoset.initial_index.value, u32_array(alpha),
                                               twiddles.handle if twiddles else 0))
    fold_circle_into_line = foldCircleIntoLine

    def decompose(self, eval_: SecureEvaluation):
        g = SecureColumnByCoords.uninitialized(eval_.len())
        lam = (C.c_uint32 * 4)()
        check(_L().tstwo_fri_decompose(eval_.values.ptrs(), eval_.domain.logSize(), g.ptrs(), lam))
        return SecureEvaluation(eval_.domain, g), list(lam)


def fold_line(eval_: LineEvaluation, alpha, twiddles=None) -> LineEvaluation:
    """free function of fri.ts:120-152"""
    return CudaBackend.friOps.foldLine(eval_, alpha, twiddles)


def fold_circle_into_line(dst: LineEvaluation, src: SecureEvaluation, alpha, twiddles=None) -> None:
    """free function of fri.ts:162-192"""
    CudaBackend.friOps.foldCircleIntoLine(dst, src, alpha, twiddles)


def accumulate(column: SecureColumnByCoords, other: SecureColumnByCoords) -> None:
    """backend/cpu/accumulation.ts:38-49"""
    if column.len() != other.len():
        raise ValueError("column length mismatch")
    check(_L().tstwo_accumulate(column.ptrs(), other.ptrs(), column.len()))


def accumulate_quotients(domain: CircleDomain, columns: Sequence[CudaCircleEvaluation], random_coeff, sample_batches,
                         log_blowup_factor: int = 1, ts_denominators: bool = True) -> SecureEvaluation:
    """backend/cpu/quotients.ts:52-75.  sample_batches: [(point=(x[4], y[4]), [(column_index, value[4]), ...]), ...]
    ts_denominators: True = bit-exact with the TS CpuBackend (denominators from the M31 parts of the point only, quotients.ts:
    160-178); False = Stwo's CM31 denominators, the only ones for which the result is a low-degree quotient."""
    pts, offs, ci, vals = [], [0], [], []
    for point, entries in sample_batches:
        pts += list(point[0]) + list(point[1])
        for col, v in entries:
            ci.append(int(col)); vals += list(v)
        offs.append(len(ci))
    out = SecureColumnByCoords.uninitialized(domain.size())
    check(_L().tstwo_accumulate_quotients_ex(domain.logSize(), domain.halfCoset.initial_index.value,
                                             dptr_array([c.values.ptr for c in columns]), len(columns), u32_array(random_coeff),
                                             len(sample_batches), u32_array(pts), (C.c_int * len(offs))(*offs),
                                             (C.c_int * max(1, len(ci)))(*ci), u32_array(vals), out.ptrs(),
                                             FLAG_QUIRKS if ts_denominators else 0))
    return SecureEvaluation(domain, out)


class CudaBackend:
    """Backend (backend/index.ts:20-41) + PolyOps batch entry points (poly/circle/ops.ts:20-101)."""

    name = "CudaBackend"
    device = 0
    # reproduce CpuBackend's log_size==3 swap; False = Stwo-correct.  Env TSTWO_QUIRKS=0 / 1 sets the default.
    quirks = os.environ.get("TSTWO_QUIRKS", "1") not in ("0", "false", "False", "")
    merkleOps = CudaMerkleOps()
    friOps = CudaFriOps()

    @staticmethod
    def createBaseFieldColumn(data) -> CudaColumn: return CudaColumn.from_cpu(data)

    @staticmethod
    def createSecureFieldColumn(data) -> SecureColumnByCoords:
        a = np.asarray(data, dtype=np.uint32).reshape(-1, 4)
        return SecureColumnByCoords.from_cpu([a[:, k] for k in range(4)])

    @staticmethod
    def bitReverseColumn(col: CudaColumn) -> None:
        n = col.len()
        if n == 0 or n & (n - 1):
            raise ValueError("length is not power of two")      # backend/cpu/index.ts:65
        check(_L().tstwo_bit_reverse(col.ptr, n))

    @staticmethod
    def precomputeTwiddles(coset: Coset) -> TwiddleTree:
        h = C.c_uint64()
        check(_L().tstwo_twiddles_precompute(coset.initial_index.value, coset.log_size, C.byref(h)))
        return TwiddleTree(coset, h.value)

    @staticmethod
    def interpolateColumns(columns: Sequence[CudaCircleEvaluation], twiddles: TwiddleTree) -> List[CudaCirclePoly]:
        """poly/circle/ops.ts:73-82; one batched launch per distinct domain."""
        out: List[Optional[CudaCirclePoly]] = [None] * len(columns)
        groups = {}
        for i, ev in enumerate(columns):
            groups.setdefault(ev.domain, []).append(i)
        for dom, idxs in groups.items():
            block = CudaColumn.uninitialized(dom.size() * len(idxs))   # one block: evenly spaced columns (bulk-copy passes)
            dst = [block.slice(k * dom.size(), dom.size()) for k in range(len(idxs))]
            try:
                check(_L().tstwo_interpolate_columns(dptr_array([columns[i].values.ptr for i in idxs]),
                                                     dptr_array([d.ptr for d in dst]), len(idxs), dom.logSize(),
                                                     dom.halfCoset.initial_index.value, twiddles.handle, _flags()))
            except TwiddleMismatch:
                raise TwiddleMismatch(-3, "twiddle tree mismatch")
            for i, d in zip(idxs, dst):
                out[i] = CudaCirclePoly(d)
        return out   # type: ignore

    @staticmethod
    def evaluatePolynomialsOnDomain(polys: Sequence[CudaCirclePoly], domain: CircleDomain, twiddles: TwiddleTree):
        out: List[Optional[CudaCircleEvaluation]] = [None] * len(polys)
        groups = {}
        for i, p in enumerate(polys):
            groups.setdefault(p.logSize(), []).append(i)
        for plog, idxs in groups.items():
            if plog > domain.logSize():
                raise ValueError("log size too small")
            block = CudaColumn.uninitialized(domain.size() * len(idxs))
            dst = [block.slice(k * domain.size(), domain.size()) for k in range(len(idxs))]
            check(_L().tstwo_evaluate_columns(dptr_array([polys[i].coeffs.ptr for i in idxs]), plog,
                                              dptr_array([d.ptr for d in dst]), len(idxs), domain.logSize(),
                                              domain.halfCoset.initial_index.value, twiddles.handle, _flags()))
            for i, d in zip(idxs, dst):
                out[i] = CudaCircleEvaluation(domain, d)
        return out

    @staticmethod
    def evaluatePolynomials(polys: Sequence[CudaCirclePoly], logBlowupFactor: int, twiddles: TwiddleTree):
        """poly/circle/ops.ts:89-101: each poly on CanonicCoset(logSize + logBlowup).circleDomain()."""
        out: List[Optional[CudaCircleEvaluation]] = [None] * len(polys)
        groups = {}
        for i, p in enumerate(polys):
            groups.setdefault(p.logSize(), []).append(i)
        for plog, idxs in groups.items():
            dom = CanonicCoset.new(plog + logBlowupFactor).circleDomain()
            evs = CudaBackend.evaluatePolynomialsOnDomain([polys[i] for i in idxs], dom, twiddles)
            for i, e in zip(idxs, evs):
                out[i] = e
        return out


class CommitmentTreeProver:
    """Stwo's CommitmentTreeProver::new (Rust text quoted in the reference at pcs/prover.ts:215-235):
    polynomials -> evaluations on the blown-up canonic domain -> MerkleProver.commit.  `from_trace` is the fused
    metric path (tstwo_lde_commit): interpolate + evaluate + commit on one contiguous column block."""

    def __init__(self, polynomials, evaluations, commitment: MerkleProver):
        self.polynomials = polynomials
        self.evaluations = evaluations
        self.commitment = commitment

    @staticmethod
    def new(polynomials: Sequence[CudaCirclePoly], log_blowup_factor: int, twiddles: TwiddleTree, channel=None) -> "CommitmentTreeProver":
        """`channel` (a Blake2sChannel) receives the root as in the Rust text (`MC::mix_root(channel, tree.root())`)."""
        evals = CudaBackend.evaluatePolynomials(polynomials, log_blowup_factor, twiddles)
        commitment = MerkleProver.commit(CudaBackend.merkleOps, [e.values for e in evals])
        if channel is not None:
            Blake2sMerkleChannel.mix_root(channel, commitment.root())
        return CommitmentTreeProver(list(polynomials), evals, commitment)

    def decommit(self, queries):
        """pcs/prover.ts:237-251 (Rust text): queried values and decommitment of this tree for {log size: positions}."""
        return self.commitment.decommit(queries, [e.values for e in self.evaluations])

    @staticmethod
    def from_trace(trace: CudaColumn, n_cols: int, log_size: int, log_blowup_factor: int, twiddles: TwiddleTree,
                   keep_coeffs: bool = True, channel=None) -> "CommitmentTreeProver":
        ext = log_size + log_blowup_factor
        n, ne = 1 << log_size, 1 << ext
        if trace.len() != n_cols * n:
            raise ValueError("trace block size mismatch")
        coeffs = CudaColumn.uninitialized(n_cols * n) if keep_coeffs else None
        evals = CudaColumn.uninitialized(n_cols * ne)
        layers = CudaColumn.uninitialized(8 * ((2 << ext) - 1))
        root = (C.c_uint8 * 32)()
        check(_L().tstwo_lde_commit(trace.ptr, n_cols, log_size, log_blowup_factor, twiddles.handle,
                                    coeffs.ptr if coeffs else 0, evals.ptr, layers.ptr, root, _flags()))
        dom = CanonicCoset.new(ext).circleDomain()
        polys = [CudaCirclePoly(coeffs.slice(c * n, n)) for c in range(n_cols)] if coeffs else []
        evs = [CudaCircleEvaluation(dom, evals.slice(c * ne, ne)) for c in range(n_cols)]
        mp = MerkleProver([layers.slice(8 * ((1 << k) - 1), 8 << k) for k in range(ext + 1)], bytes(root))
        if channel is not None:
            Blake2sMerkleChannel.mix_root(channel, mp.root())
        return CommitmentTreeProver(polys, evs, mp)

    def root(self) -> bytes: return self.commitment.root()


class FriConfig:
    """fri.ts:35-70 (log_last_layer_degree_bound, log_blowup_factor, n_queries) with the reference's range checks."""
    LOG_MIN_LAST_LAYER_DEGREE_BOUND, LOG_MAX_LAST_LAYER_DEGREE_BOUND = 0, 10
    LOG_MIN_BLOWUP_FACTOR, LOG_MAX_BLOWUP_FACTOR = 1, 16

    def __init__(self, log_last_layer_degree_bound: int, log_blowup_factor: int, n_queries: int):
        if not self.LOG_MIN_LAST_LAYER_DEGREE_BOUND <= log_last_layer_degree_bound <= self.LOG_MAX_LAST_LAYER_DEGREE_BOUND:
            raise ValueError("invalid log_last_layer_degree_bound")
        if not self.LOG_MIN_BLOWUP_FACTOR <= log_blowup_factor <= self.LOG_MAX_BLOWUP_FACTOR:
            raise ValueError("invalid log_blowup_factor")
        self.log_last_layer_degree_bound = log_last_layer_degree_bound
        self.log_blowup_factor = log_blowup_factor
        self.n_queries = n_queries

    def last_layer_domain_size(self) -> int:
        return 1 << (self.log_last_layer_degree_bound + self.log_blowup_factor)


class PcsConfig:
    """pcs/index.ts: pow_bits + FriConfig."""

    def __init__(self, pow_bits: int = 5, fri_config: Optional[FriConfig] = None):
        self.pow_bits = pow_bits
        self.fri_config = fri_config or FriConfig(0, 1, 3)


class TreeSubspan:
    def __init__(self, tree_index: int, col_start: int, col_end: int):
        self.tree_index, self.col_start, self.col_end = tree_index, col_start, col_end

    def __eq__(self, o): return (self.tree_index, self.col_start, self.col_end) == (o.tree_index, o.col_start, o.col_end)
    def __repr__(self): return f"TreeSubspan({self.tree_index}, {self.col_start}, {self.col_end})"


class TreeBuilder:
    """pcs/prover.ts:170-213 (Rust text): collects the polynomials of one commitment tree."""

    def __init__(self, scheme: "CommitmentSchemeProver"):
        self.tree_index = len(scheme.trees)
        self.commitment_scheme = scheme
        self.polys: List[CudaCirclePoly] = []

    def extend_evals(self, columns: Sequence[CudaCircleEvaluation]) -> TreeSubspan:
        return self.extend_polys(CudaBackend.interpolateColumns(list(columns), self.commitment_scheme.twiddles))

    def extend_polys(self, columns: Sequence[CudaCirclePoly]) -> TreeSubspan:
        start = len(self.polys)
        self.polys.extend(columns)
        return TreeSubspan(self.tree_index, start, len(self.polys))

    def commit(self, channel) -> None:
        self.commitment_scheme._commit(self.polys, channel)


class CommitmentSchemeProver:
    """pcs/prover.ts:28-84 (Rust text; never ported by the reference): the trees committed so far, one
    CommitmentTreeProver each, all on the same twiddle tree; every commit mixes its root into the channel."""

    def __init__(self, config: PcsConfig, twiddles: TwiddleTree):
        self.trees: List[CommitmentTreeProver] = []
        self.config = config
        self.twiddles = twiddles

    def _commit(self, polynomials: Sequence[CudaCirclePoly], channel) -> None:
        self.trees.append(CommitmentTreeProver.new(polynomials, self.config.fri_config.log_blowup_factor, self.twiddles, channel))

    def tree_builder(self) -> TreeBuilder: return TreeBuilder(self)
    def roots(self) -> List[bytes]: return [t.commitment.root() for t in self.trees]
    def polynomials(self): return [list(t.polynomials) for t in self.trees]
    def evaluations(self): return [list(t.evaluations) for t in self.trees]

    def decommit(self, queries):
        """The Merkle part of prove_values (pcs/prover.ts:139-146): per tree (queried values, decommitment)."""
        return [t.decommit(queries) for t in self.trees]

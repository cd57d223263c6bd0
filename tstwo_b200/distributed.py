"""Multi-GPU LDE + Merkle commit: one process per GPU over torch.distributed (BASELINE config 3).

Data flow for G ranks, n_cols columns of 2^log_size, blow-up 2^b (ext = log_size + b):
  1. every rank interpolates + extends its n_cols/G columns (column-sharded LDE); the evaluations are written
     grouped by destination rank, [G][n_cols/G][2^ext / G] (tstwo_lde_columns_sharded);
  2. ONE all-to-all re-shards them by row range: rank d receives [G][n_cols/G][rows] = [n_cols][rows], i.e. all
     columns of its contiguous row range -- a complete Merkle subtree of height ext - log2 G;
  3. every rank hashes its leaves and builds its subtree (tstwo_merkle_commit);
  4. the G subtree roots are all-gathered; rank 0 finishes the top log2 G layers (at most G-1 compressions).
The root equals the single-GPU root bit for bit (tests/test_distributed_cpu.py, tests/test_gpu_pipeline.py).

`ops` abstracts the device work so the host protocol can be exercised on CPU (gloo) with another provider; the
product provider is CudaShardOps below (C ABI only, no fallback).
"""
from __future__ import annotations

import ctypes as C
from typing import List

import torch
import torch.distributed as dist


class CudaShardOps:
    """Device work of the sharded commit through libtstwo_cuda.so on torch CUDA tensors."""

    def __init__(self, device_index: int, stream=None):
        from . import _lib
        self._lib_mod = _lib
        self.L = _lib.init(device_index)
        self.dev = torch.device("cuda", device_index)
        self._stream = stream
        if stream is not None:
            _lib.check(self.L.tstwo_set_stream(C.c_void_p(stream.cuda_stream)))

    # torch.distributed collectives run on torch's CURRENT stream, the library's kernels on the library stream: the same
    # stream only if the caller passed `stream` and made it current.  Otherwise order them explicitly, so that a collective
    # neither overtakes the kernels before it nor is overtaken by the ones after it (with stream=None the library uses its own
    # non-blocking stream, which torch cannot see: drain it).
    def _same_stream(self) -> bool:
        return self._stream is not None and torch.cuda.current_stream(self.dev).cuda_stream == self._stream.cuda_stream

    def before_collective(self) -> None:
        if self._same_stream():
            return
        if self._stream is not None:
            torch.cuda.current_stream(self.dev).wait_stream(self._stream)
        else:
            self._lib_mod.check(self.L.tstwo_sync())

    def after_collective(self) -> None:
        if self._same_stream():
            return
        if self._stream is not None:
            self._stream.wait_stream(torch.cuda.current_stream(self.dev))
        else:
            torch.cuda.current_stream(self.dev).synchronize()

    def empty(self, n_words: int) -> torch.Tensor:
        return torch.empty(n_words, dtype=torch.int32, device=self.dev)

    def twiddles(self, ext: int) -> int:
        h = C.c_uint64()
        self._lib_mod.check(self.L.tstwo_twiddles_precompute(1 << (31 - (ext - 1) - 2), ext - 1, C.byref(h)))
        return h.value

    def lde_sharded(self, trace: torch.Tensor, n_cols: int, log_size: int, log_blowup: int, tw: int, coeffs, evals: torch.Tensor,
                    n_shards: int) -> None:
        self._lib_mod.check(self.L.tstwo_lde_columns_sharded(trace.data_ptr(), n_cols, log_size, log_blowup, tw,
                                                             coeffs.data_ptr() if coeffs is not None else 0, evals.data_ptr(),
                                                             n_shards, 0))

    def merkle_commit(self, block: torch.Tensor, n_cols: int, log_rows: int, layers: torch.Tensor) -> None:
        rows = 1 << log_rows
        ptrs = self._lib_mod.dptr_array([block.data_ptr() + 4 * rows * c for c in range(n_cols)])
        logs = (C.c_int * n_cols)(*([log_rows] * n_cols))
        self._lib_mod.check(self.L.tstwo_merkle_commit(ptrs, logs, n_cols, layers.data_ptr(), None))

    def commit_layer(self, log: int, prev: torch.Tensor, out: torch.Tensor) -> None:
        self._lib_mod.check(self.L.tstwo_merkle_commit_layer(log, prev.data_ptr(), None, 0, out.data_ptr()))

    def to_host_bytes(self, t: torch.Tensor) -> bytes:
        return t.cpu().numpy().tobytes()


def top_layers(ops, roots: torch.Tensor, world: int) -> List[torch.Tensor]:
    """Finish the tree above the subtree roots: roots = [world][8 words]; returns layers log2(world)-1 .. 0."""
    out = []
    prev = roots
    for k in range(world.bit_length() - 2, -1, -1):
        cur = ops.empty(8 << k)
        ops.commit_layer(k, prev, cur)
        out.append(cur)
        prev = cur
    return out


def owned_columns(rank: int, world: int, n_cols: int, ownership: str = "contiguous"):
    """The trace columns of a rank, in the order of its local block: [(first column, count), ...].
    "contiguous": columns [rank c, (rank + 1) c), c = n_cols / world.  "halves": the first c / 2 local columns are columns
    rank c / 2 .. of the trace's first half, the others the same range of its second half -- after half of every rank's work the
    receivers hold columns [0, n_cols / 2) of their rows, a prefix of every leaf's message, and can start hashing it."""
    lc = n_cols // world
    if ownership == "contiguous":
        return [(rank * lc, lc)]
    if ownership != "halves":
        raise ValueError("ownership must be 'contiguous' or 'halves'")
    if lc % 2:
        raise ValueError("column halves need an even number of columns per rank")
    h = lc // 2
    return [(rank * h, h), (n_cols // 2 + rank * h, h)]


class ShardedCommit:
    """Buffers + one `step()` of the multi-rank pipeline.  trace: this rank's [n_cols/G][2^log_size] block."""

    def __init__(self, ops, n_cols: int, log_size: int, log_blowup: int, keep_coeffs: bool = True):
        self.ops = ops
        self.world = dist.get_world_size()
        self.rank = dist.get_rank()
        G = self.world
        if G & (G - 1) or n_cols % G:
            raise ValueError("world size must be a power of two dividing n_cols")
        self.n_cols, self.lc = n_cols, n_cols // G
        self.log_size, self.blow, self.ext = log_size, log_blowup, log_size + log_blowup
        self.sub = self.ext - (G.bit_length() - 1)               # log2 rows per rank
        if self.sub < 12:
            raise ValueError("row shards must hold at least 2^12 rows")
        rows = 1 << self.sub
        self.tw = ops.twiddles(self.ext)
        self.coeffs = ops.empty(self.lc << log_size) if keep_coeffs else None
        self.send = ops.empty(self.lc << self.ext)               # [G][lc][rows]
        self.recv = ops.empty(n_cols * rows)                     # [G][lc][rows] == [n_cols][rows]
        self.layers = ops.empty(8 * ((2 << self.sub) - 1))       # this rank's subtree, root first
        self.roots = ops.empty(8 * G)
        self.top: List[torch.Tensor] = []

    def step(self, trace: torch.Tensor) -> bytes | None:
        ops = self.ops
        before = getattr(ops, "before_collective", lambda: None)
        after = getattr(ops, "after_collective", lambda: None)
        ops.lde_sharded(trace, self.lc, self.log_size, self.blow, self.tw, self.coeffs, self.send, self.world)
        before(); dist.all_to_all_single(self.recv, self.send); after()
        ops.merkle_commit(self.recv, self.n_cols, self.sub, self.layers)
        before(); dist.all_gather_into_tensor(self.roots, self.layers[:8].contiguous()); after()
        if self.rank == 0:
            self.top = top_layers(ops, self.roots, self.world)
            return ops.to_host_bytes(self.top[-1] if self.top else self.roots)[:32]
        return None


class PeerShardedCommit:
    """The same pipeline with the exchange FUSED into the last FFT pass (no send buffer, no all-to-all): every rank maps
    the leaf buffers of all ranks (CUDA IPC over NVLink) and its last evaluate pass stores each row range straight into
    the owner's [n_cols][rows] buffer (tstwo_lde_columns_to_peers) while the butterflies of the next tiles run.
      1. LDE of the n_cols/G local columns, evaluations stored into the G leaf buffers at column rank * n_cols/G + c;
      2. one 4-byte all-reduce orders the step: when it completes on a rank, every rank's stores have been issued and
         completed (stream order + kernel completion), so the rank may hash its leaves;
      3. subtree per rank, all-gather of the G roots (which also keeps a fast rank from overwriting a leaf buffer that
         is still being hashed in the next step), top log2 G layers on rank 0.
    CUDA-only (needs peer access); the protocol above the device work is the one tests/test_distributed_cpu.py covers."""

    # word offsets inside the 4 KiB flag block that follows every rank's leaf buffer (same CUDA IPC mapping)
    _ARRIVE, _HASHED, _ROOTF, _HALF, _ROOTS, FLAG_BYTES = 0, 64, 128, 192, 256, 4096

    def __init__(self, device_index: int, stream, n_cols: int, log_size: int, log_blowup: int, keep_coeffs: bool = False,
                 overlap: tuple | None = None, fence: str = "flags", ownership: str = "contiguous"):
        """overlap = (col_groups, peer_sms): run the NVLink-bound last pass of each column group on a side stream on peer_sms
        SMs while the others transform the next group (tstwo_lde_columns_to_peers_overlapped); None = one group.
        fence = "flags": the ranks order a step with release/acquire flags in each other's memory (tstwo_peer_signal / _wait /
        _put: no NCCL kernel in the step, no host synchronisation); "nccl": the 4-byte all-reduce + 32-byte all-gather of round 1.
        ownership = "halves" (owned_columns; needs fence = "flags" and an overlap with an even number of groups): `trace` holds
        the rank's columns of the first half of the trace, then those of the second half; the receivers hash the first half of
        every leaf (tstwo_merkle_leaf_absorb) beside the last group's storing pass instead of after it.  Same root."""
        from . import _lib
        if fence not in ("flags", "nccl"):
            raise ValueError("fence must be 'flags' or 'nccl'")
        if ownership not in ("contiguous", "halves"):
            raise ValueError("ownership must be 'contiguous' or 'halves'")
        if ownership == "halves" and not self.halves_supported(dist.get_world_size(), n_cols, log_size, overlap, fence):
            raise ValueError("column halves need fence='flags', an overlap (even number of groups dividing the rank's columns, "
                             "peer_sms > 0), n_cols a multiple of 32 and log_size >= 15")
        self.ownership = ownership
        self.fence = fence
        self._step = 0
        self.overlap = overlap
        self._lib_mod = _lib
        self.L = _lib.init(device_index)
        chk = _lib.check
        self.dev = torch.device("cuda", device_index)
        self._stream = stream
        if stream is not None:
            chk(self.L.tstwo_set_stream(C.c_void_p(stream.cuda_stream)))
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        G = self.world
        if G & (G - 1) or n_cols % G:
            raise ValueError("world size must be a power of two dividing n_cols")
        self.n_cols, self.lc = n_cols, n_cols // G
        self.log_size, self.blow, self.ext = log_size, log_blowup, log_size + log_blowup
        self.sub = self.ext - (G.bit_length() - 1)
        if self.sub < 14:
            raise ValueError("row shards must hold at least 2^14 rows")
        rows = 1 << self.sub
        self.ops = CudaShardOps(device_index)      # merkle / top layers
        self.tw = self.ops.twiddles(self.ext)
        self.coeffs = self.ops.empty(self.lc << log_size) if keep_coeffs else None
        self.work = self.ops.empty(self.lc << self.ext)
        self.layers = self.ops.empty(8 * ((2 << self.sub) - 1))
        self.roots = self.ops.empty(8 * G)
        self.flag = torch.zeros(1, dtype=torch.int32, device=self.dev)
        self.top: List[torch.Tensor] = []
        # leaf buffer of this rank: its own cudaMalloc (IPC exports whole allocations), exported to every peer
        p = C.c_uint64()
        self.leaf_bytes = 4 * n_cols * rows
        chk(self.L.tstwo_malloc(self.leaf_bytes + self.FLAG_BYTES, C.byref(p)))
        self.recv_ptr = p.value
        chk(self.L.tstwo_memset_zero(self.recv_ptr + self.leaf_bytes, self.FLAG_BYTES))
        chk(self.L.tstwo_sync())
        handle = (C.c_uint8 * 64)()
        chk(self.L.tstwo_ipc_export(self.recv_ptr, handle))
        mine = torch.tensor(list(bytes(handle)), dtype=torch.uint8, device=self.dev)
        allh = torch.empty(64 * G, dtype=torch.uint8, device=self.dev)
        dist.all_gather_into_tensor(allh, mine)
        allh = allh.cpu().numpy()
        self.bases, self._imported = [], []
        for r in range(G):
            if r == self.rank:
                self.bases.append(self.recv_ptr)
                continue
            q = C.c_uint64()
            h = (C.c_uint8 * 64)(*allh[64 * r:64 * (r + 1)].tolist())
            chk(self.L.tstwo_ipc_import(h, C.byref(q)))
            self.bases.append(q.value)
            self._imported.append(q.value)
        self._bases_arr = _lib.dptr_array(self.bases)
        # per-step argument arrays, built once (the step is a few ms: host work per step counts)
        G, r = self.world, self.rank
        fb = [b + self.leaf_bytes for b in self.bases]
        self._flag_local = fb[r]
        self._arrive_addrs = _lib.dptr_array([f + 4 * (self._ARRIVE + r) for f in fb])
        self._hashed_addrs = _lib.dptr_array([f + 4 * (self._HASHED + r) for f in fb])
        self._half_addrs = _lib.dptr_array([f + 4 * (self._HALF + r) for f in fb])
        self._root_dst = fb[0] + 4 * (self._ROOTS + 8 * r)
        self._root_flag = fb[0] + 4 * (self._ROOTF + r)
        self._leaf_ptrs = _lib.dptr_array([self.recv_ptr + 4 * rows * c for c in range(self.n_cols)])
        self._leaf_logs = (C.c_int * self.n_cols)(*([self.sub] * self.n_cols))
        self._top = [self.ops.empty(8 << k) for k in range(G.bit_length() - 2, -1, -1)] if r == 0 else []
        import os
        self._timeout_ms = int(os.environ.get("TSTWO_PEER_TIMEOUT_MS", "10000"))
        dist.barrier()                             # every rank's flags are zero and mapped before anyone signals

    @staticmethod
    def halves_supported(world: int, n_cols: int, log_size: int, overlap, fence: str = "flags") -> bool:
        """Whether ownership = "halves" applies to this shape: the two leaf chunks must be whole BLAKE2s blocks (n_cols / 2 a
        multiple of 16), the rank's columns split into an even number of equal groups, and the LDE must have a storing pass to put
        on a side stream (log_size >= 15)."""
        if world < 1 or n_cols % world:
            return False
        lc = n_cols // world
        return (fence == "flags" and bool(overlap) and 2 <= overlap[0] <= 32 and overlap[0] % 2 == 0 and overlap[1] >= 1 and lc % overlap[0] == 0
                and n_cols % 32 == 0 and log_size >= 15)

    def _before_stores(self):
        """My stores go into every rank's leaf buffer: their hashing of the previous step must be over (flags: every rank
        publishes hashed[rank] = step after its tree; nccl: the all-gather at the end of the previous step said so)."""
        if self.fence == "flags" and self._step > 0:
            self._lib_mod.check(self.L.tstwo_peer_wait(self._flag_local + 4 * self._HASHED, self.world, self._step, self._timeout_ms))

    def step(self, trace: torch.Tensor, read_root: bool = True):
        """One LDE + commit.  read_root = False leaves the root on the device (rank 0: root_bytes()), so that the host can enqueue
        the next step while this one runs."""
        self._before_stores()
        self._lde(trace)
        return self._finish(read_root=read_root, split=self.ownership == "halves")

    def _lde(self, trace: torch.Tensor):
        """The rank's LDE with its stores into the peers' leaf buffers (halves: also the half / arrive flags, side stream left
        unjoined -- _finish hashes beside it)."""
        chk = self._lib_mod.check
        groups, peer_sms = self.overlap if self.overlap else (1, 0)
        cf = self.coeffs.data_ptr() if self.coeffs is not None else 0
        if self.ownership == "halves":
            h = self.lc // 2
            chk(self.L.tstwo_lde_columns_to_peers_halves(trace.data_ptr(), self.lc, self.log_size, self.blow, self.tw, cf,
                                                         self.work.data_ptr(), self._bases_arr, self.world, self.rank * h,
                                                         self.n_cols // 2 + self.rank * h, 0, groups, peer_sms,
                                                         self._half_addrs, self._arrive_addrs, self.world, self._step + 1))
        else:
            chk(self.L.tstwo_lde_columns_to_peers_overlapped(trace.data_ptr(), self.lc, self.log_size, self.blow, self.tw, cf,
                                                             self.work.data_ptr(), self._bases_arr, self.world, self.rank * self.lc, 0,
                                                             groups, peer_sms))

    def _dst_col(self, local_col: int) -> int:
        """Trace column of local column `local_col` (owned_columns)."""
        for first, count in owned_columns(self.rank, self.world, self.n_cols, self.ownership):
            if local_col < count:
                return first + local_col
            local_col -= count
        raise IndexError(local_col)

    def _chunks(self, chunks: int) -> int:
        """Largest chunk count <= chunks that divides the rank's columns and keeps every chunk inside one ownership segment."""
        seg = self.lc // len(owned_columns(self.rank, self.world, self.n_cols, self.ownership))
        chunks = max(1, min(chunks, self.lc))
        while self.lc % chunks or seg % (self.lc // chunks):
            chunks -= 1
        return chunks

    def root_bytes(self) -> bytes | None:
        """The last step's root (rank 0; synchronises)."""
        if self.rank != 0:
            return None
        t = self._top[-1] if self._top else None
        if t is None:                              # one rank: the subtree root is the root
            return self.ops.to_host_bytes(self.layers[:8])[:32]
        return self.ops.to_host_bytes(t)[:32]

    def step_phases(self, trace: torch.Tensor):
        """step() with a CUDA event at every phase boundary (for bench.py's per-rank breakdown): returns the 5 events around
        [LDE + peer stores | fence (all-reduce) | leaves + subtree | root gather + top layers]."""
        st = self._stream if self._stream is not None else torch.cuda.current_stream()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
        self._before_stores()
        ev[0].record(st)
        self._lde(trace)
        if self.ownership != "halves":         # halves: the phase ends where the side stream is joined (inside _finish)
            ev[1].record(st)
        self._finish(ev, st, split=self.ownership == "halves")
        return ev

    def _absorb_first_half(self, s: int):
        """Once every rank has published half[rank] = s: hash columns [0, n_cols / 2) of my rows (chaining values in the leaf slots)."""
        chk, L = self._lib_mod.check, self.L
        chk(L.tstwo_peer_wait(self._flag_local + 4 * self._HALF, self.world, s, self._timeout_ms))
        chk(L.tstwo_merkle_leaf_absorb(self.recv_ptr, 1 << self.sub, 0, self.n_cols // 2, self.n_cols, self.sub, self.layers.data_ptr()))

    def _finish(self, ev=None, st=None, read_root: bool = True, split=False):
        """split = True: _lde ran in halves mode (flags published by its side stream, side stream not yet joined).
        split = "absorbed": the caller has hashed the first half of the columns already (_absorb_first_half) and nothing runs on a
        side stream; arrive[rank] is published here."""
        chk = self._lib_mod.check
        L, G = self.L, self.world
        self._step += 1
        s = self._step
        if split:
            # the first half of every leaf is hashed beside my last storing pass (still running on peer_sms SMs of the side stream,
            # which has published half[rank] / arrive[rank] = s to every rank), the second once everything has arrived
            rows, half = 1 << self.sub, self.n_cols // 2
            if split == "absorbed":
                chk(L.tstwo_peer_signal(self._arrive_addrs, G, s))
            else:
                self._absorb_first_half(s)
                chk(L.tstwo_peer_join())
            if ev: ev[1].record(st)
            chk(L.tstwo_peer_wait(self._flag_local + 4 * self._ARRIVE, G, s, self._timeout_ms))
            if ev: ev[2].record(st)
            chk(L.tstwo_merkle_leaf_absorb(self.recv_ptr + 4 * rows * half, rows, half, half, self.n_cols, self.sub, self.layers.data_ptr()))
            chk(L.tstwo_merkle_commit_nodes(self.sub, self.layers.data_ptr(), None))
            if ev: ev[3].record(st)
        else:
            if self.fence == "flags":
                # "my stores are complete" into every rank's arrive[rank]; then wait until every rank has said so here
                chk(L.tstwo_peer_signal(self._arrive_addrs, G, s))
                chk(L.tstwo_peer_wait(self._flag_local + 4 * self._ARRIVE, G, s, self._timeout_ms))
            else:
                self._order_before_collective()
                dist.all_reduce(self.flag)             # every rank's stores into my leaf buffer are complete after this
                self._order_after_collective()
            if ev: ev[2].record(st)
            chk(L.tstwo_merkle_commit(self._leaf_ptrs, self._leaf_logs, self.n_cols, self.layers.data_ptr(), None))
            if ev: ev[3].record(st)
        if self.fence == "flags":
            chk(L.tstwo_peer_signal(self._hashed_addrs, G, s))                     # my leaf buffer may be overwritten
            chk(L.tstwo_peer_put(self._root_dst, self.layers.data_ptr(), 8, self._root_flag, s))
            if self.rank == 0:
                chk(L.tstwo_peer_wait(self._flag_local + 4 * self._ROOTF, G, s, self._timeout_ms))
                prev = self._flag_local + 4 * self._ROOTS
                for i, k in enumerate(range(G.bit_length() - 2, -1, -1)):
                    chk(L.tstwo_merkle_commit_layer(k, prev, None, 0, self._top[i].data_ptr()))
                    prev = self._top[i].data_ptr()
                self.top = self._top
        else:
            self._order_before_collective()
            dist.all_gather_into_tensor(self.roots, self.layers[:8].contiguous())
            self._order_after_collective()
            if self.rank == 0:
                self.top = top_layers(self.ops, self.roots, self.world)
                self._top = self.top
        out = self.root_bytes() if (read_root and self.rank == 0) else None
        if ev: ev[4].record(st)
        return out

    # torch.distributed collectives run on torch's CURRENT stream, the library's kernels on the library stream.  They are the
    # same stream only when the caller passed `stream` and made it current (bench.py does).  Otherwise the two are ordered
    # here with events, so that a collective never overtakes the stores before it nor the hashing after it.
    def _lib_stream_is_current(self) -> bool:
        return self._stream is not None and torch.cuda.current_stream(self.dev).cuda_stream == self._stream.cuda_stream

    def _order_before_collective(self):
        if self._lib_stream_is_current():
            return
        if self._stream is not None:
            torch.cuda.current_stream(self.dev).wait_stream(self._stream)
        else:
            self._lib_mod.check(self.L.tstwo_sync())       # library-owned stream: nothing to wait on from torch, drain it

    def _order_after_collective(self):
        if self._lib_stream_is_current():
            return
        if self._stream is not None:
            self._stream.wait_stream(torch.cuda.current_stream(self.dev))
        else:
            torch.cuda.current_stream(self.dev).synchronize()

    def step_pipelined(self, trace: torch.Tensor, chunks: int = 4):
        """The exchange on the copy engines instead of inside the last FFT pass: the local columns are transformed in `chunks`
        groups, each written grouped by destination rank (tstwo_lde_columns_sharded) and pushed to the owners' leaf buffers by
        device-to-device copies on side streams (tstwo_scatter_shards) while the SMs transform the next group.  Only the last
        group's copies are exposed; the SMs never wait on NVLink."""
        chk = self._lib_mod.check
        self._before_stores()
        chunks = self._chunks(chunks)
        per, n, rows = self.lc // chunks, 1 << self.log_size, 1 << self.sub
        for k in range(chunks):
            work = self.work.data_ptr() + 4 * ((k * per) << self.ext)
            chk(self.L.tstwo_lde_columns_sharded(trace.data_ptr() + 4 * n * k * per, per, self.log_size, self.blow, self.tw,
                                                 self.coeffs.data_ptr() + 4 * n * k * per if self.coeffs is not None else 0,
                                                 work, self.world, 0))
            chk(self.L.tstwo_scatter_shards(work, per, rows, self._bases_arr, self.world, self._dst_col(k * per)))
        chk(self.L.tstwo_scatter_join())
        return self._finish()

    def step_from_host(self, host: torch.Tensor, staging: torch.Tensor, chunks: int = 4):
        """The same step with this rank's columns in pinned HOST memory (`host`, (n_cols/G, 2^log) int32): the columns cross PCIe
        in `chunks` groups on a copy stream and each group's LDE + peer stores start as soon as it has landed, so the transfer
        of the later groups overlaps the transforms and the NVLink exchange of the earlier ones."""
        chk = self._lib_mod.check
        main = self._stream if self._stream is not None else torch.cuda.current_stream()
        if not hasattr(self, "_copy_stream"):
            self._copy_stream = torch.cuda.Stream(device=self.dev)
        cs = self._copy_stream
        chunks = self._chunks(chunks)
        per, n = self.lc // chunks, 1 << self.log_size
        early = self.ownership == "halves" and self.fence == "flags" and chunks >= 2
        self._before_stores()
        cs.wait_stream(main)                       # the staging buffer is free once the previous step's transforms are done
        for k in range(chunks):
            with torch.cuda.stream(cs):
                staging[k * per:(k + 1) * per].copy_(host[k * per:(k + 1) * per], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(cs)
            main.wait_event(ev)
            chk(self.L.tstwo_lde_columns_to_peers(staging.data_ptr() + 4 * n * k * per, per, self.log_size, self.blow, self.tw,
                                                  self.coeffs.data_ptr() + 4 * n * k * per if self.coeffs is not None else 0,
                                                  self.work.data_ptr() + 4 * ((k * per) << self.ext),
                                                  self._bases_arr, self.world, self._dst_col(k * per), 0))
            if early and (k + 1) * per == self.lc // 2:
                # halves: my share of the trace's first half is stored.  Once every rank's is, hash that half of every leaf -- while
                # the second half is still crossing PCIe, where the device has nothing else to do
                chk(self.L.tstwo_peer_signal(self._half_addrs, self.world, self._step + 1))
                self._absorb_first_half(self._step + 1)
        return self._finish(split="absorbed" if early else False)

    def close(self):
        torch.cuda.synchronize()
        dist.barrier()
        for q in self._imported:
            self.L.tstwo_ipc_close(q)
        self._imported = []
        if self.recv_ptr:
            self.L.tstwo_free(self.recv_ptr)
            self.recv_ptr = 0

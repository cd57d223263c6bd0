"""Host protocol of the multi-GPU path (tstwo_b200/distributed.py) on CPU: world_size 2 and 4 over gloo, with the
oracle standing in for the device work.  Checks the shard layout, the all-to-all re-sharding and the top-of-tree
combine against the single-process oracle root."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

P = 2147483647
LOG, BLOW, N_COLS = 13, 1, 8


class OracleShardOps:
    """CPU provider with the interface of CudaShardOps (test infrastructure only)."""

    def __init__(self):
        import oracle
        self.orc = oracle

    def empty(self, n_words):
        return torch.zeros(n_words, dtype=torch.int32)

    def twiddles(self, ext):
        init, log = self.orc.canonic_half_coset(ext)
        return self.orc.precompute_twiddles(init, log)

    def lde_sharded(self, trace, n_cols, log_size, log_blowup, tw, coeffs, evals, n_shards):
        t = trace.numpy().view(np.uint32).reshape(n_cols, 1 << log_size)
        cf, ev, _ = self.orc.lde_commit(t, log_blowup, tw[0], tw[1], want_coeffs=True, want_layers=False)
        rows = ev.shape[1] // n_shards
        out = np.stack([ev[:, s * rows:(s + 1) * rows] for s in range(n_shards)])       # [G][lc][rows]
        evals.numpy().view(np.uint32)[:] = out.reshape(-1)
        if coeffs is not None:
            coeffs.numpy().view(np.uint32)[:] = cf.reshape(-1)

    def merkle_commit(self, block, n_cols, log_rows, layers):
        b = block.numpy().view(np.uint32).reshape(n_cols, 1 << log_rows)
        lay = self.orc.merkle_commit(list(b))
        layers.numpy().view(np.uint8)[:] = np.concatenate(lay).reshape(-1)

    def commit_layer(self, log, prev, out):
        h = self.orc.merkle_commit_layer(log, prev.numpy().view(np.uint8), [])
        out.numpy().view(np.uint8)[:] = h.reshape(-1)

    def to_host_bytes(self, t):
        return t.numpy().tobytes()


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from tstwo_b200.distributed import ShardedCommit
        rng = np.random.default_rng(1234)
        full = rng.integers(0, P, (N_COLS, 1 << LOG), dtype=np.uint32)
        lc = N_COLS // world
        mine = torch.from_numpy(full[rank * lc:(rank + 1) * lc].copy().view(np.int32).reshape(-1))
        sc = ShardedCommit(OracleShardOps(), N_COLS, LOG, BLOW)
        root = sc.step(mine)
        if rank == 0:
            q.put(root)
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


@pytest.mark.parametrize("world", [2, 4])
def test_sharded_commit_matches_single_process(orc, world):
    rng = np.random.default_rng(1234)
    full = rng.integers(0, P, (N_COLS, 1 << LOG), dtype=np.uint32)
    init, log = orc.canonic_half_coset(LOG + BLOW)
    tw, itw = orc.precompute_twiddles(init, log)
    _, _, layers = orc.lde_commit(full, BLOW, tw, itw)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    root = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert root == layers[0][0].tobytes()


def test_owned_columns_partition_the_trace_and_halves_give_a_prefix():
    """distributed.owned_columns: every column belongs to exactly one rank under both ownerships; under "halves" the first segments
    of all ranks tile columns [0, n/2) -- the prefix of every leaf the receivers hash before the second half has arrived."""
    from tstwo_b200.distributed import owned_columns
    for n_cols in (32, 256):
        for world in (1, 2, 4, 8):
            for own in ("contiguous", "halves"):
                segs = [owned_columns(r, world, n_cols, own) for r in range(world)]
                cols = [c for s in segs for a, k in s for c in range(a, a + k)]
                assert sorted(cols) == list(range(n_cols))
                assert all(sum(k for _, k in s) == n_cols // world for s in segs)
            first = sorted(c for r in range(world) for a, k in owned_columns(r, world, n_cols, "halves")[:1] for c in range(a, a + k))
            assert first == list(range(n_cols // 2))
    import pytest
    with pytest.raises(ValueError):
        owned_columns(0, 8, 8, "halves")          # one column per rank cannot be split
    with pytest.raises(ValueError):
        owned_columns(0, 2, 8, "thirds")


def test_halves_ownership_reassembles_the_rows_and_hashes_in_two_chunks():
    """The data flow of PeerShardedCommit(ownership="halves") on the host: every rank stores its two column segments at their trace
    columns of the receivers' leaf buffers (first segments first), and a leaf hashed as BLAKE2s over columns [0, n/2) continued over
    [n/2, n) -- what tstwo_merkle_leaf_absorb does in two calls -- is the leaf of the whole row."""
    import hashlib
    import numpy as np
    from tstwo_b200.distributed import owned_columns
    n_cols, rows, world = 64, 8, 4
    trace = np.random.default_rng(5).integers(0, 2**31 - 1, (n_cols, rows), dtype=np.uint32)
    leaf_buffer = np.zeros_like(trace)
    arrived_after_first_phase = set()
    for phase in (0, 1):
        for r in range(world):
            segs = owned_columns(r, world, n_cols, "halves")
            local = np.concatenate([trace[a:a + c] for a, c in segs])            # the rank's local block, first-half share first
            at = sum(c for _, c in segs[:phase])
            first, count = segs[phase]
            leaf_buffer[first:first + count] = local[at:at + count]
        if phase == 0:
            arrived_after_first_phase = {c for c in range(n_cols) if leaf_buffer[c].any()}
    assert arrived_after_first_phase == set(range(n_cols // 2)), "after every rank's first segment the receivers hold a PREFIX of each row"
    assert (leaf_buffer == trace).all()
    for i in range(rows):
        h = hashlib.blake2s()
        h.update(leaf_buffer[:n_cols // 2, i].astype("<u4").tobytes())
        h.update(leaf_buffer[n_cols // 2:, i].astype("<u4").tobytes())
        assert h.digest() == hashlib.blake2s(trace[:, i].astype("<u4").tobytes()).digest()


def test_bench_make_trace_is_the_same_job_under_every_world_size_and_ownership():
    """bench.make_trace: the shards of every world size and of both column ownerships reassemble to the SAME trace (the one whose
    root bench.py prints and tests/test_gpu_config_sizes.py checks against the oracle) -- on the CPU generator here, which has its
    own random stream but the same block / seed structure."""
    import os
    import sys
    import torch
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    from tstwo_b200.distributed import owned_columns
    n_cols, n = 64, 32
    dev = torch.device("cpu")
    whole = bench.make_trace(torch, dev, 0, 1, n_cols, n)
    assert whole.shape == (n_cols, n) and int(whole.max()) < 2**31 - 1 and int(whole.min()) >= 0
    for world in (2, 4, 8):
        for own in ("contiguous", "halves"):
            got = torch.empty_like(whole)
            for r in range(world):
                shard = bench.make_trace(torch, dev, r, world, n_cols, n, own)
                at = 0
                for first, count in owned_columns(r, world, n_cols, own):
                    got[first:first + count] = shard[at:at + count]
                    at += count
            assert torch.equal(got, whole), (world, own)

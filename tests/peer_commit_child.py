"""PeerShardedCommit (one process per GPU, fused peer stores, flag fences) against the ORACLE's root, for both column ownerships
and the chunked steps; run under torchrun (tests/test_gpu_peer_commit.py does, for every world size the box offers).
usage: torchrun --nproc-per-node G tests/peer_commit_child.py n_cols log blow groups peer_sms"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

import oracle as orc
from tstwo_b200.distributed import PeerShardedCommit, owned_columns

n_cols, log, blow, groups, peer_sms = (int(x) for x in sys.argv[1:6])
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
P = 2**31 - 1
trace = np.random.default_rng(77).integers(0, P, (n_cols, 1 << log), dtype=np.uint32)      # the same on every rank
want = None
if rank == 0:
    ri, rl = orc.canonic_half_coset(log + blow)
    otw, oitw = orc.precompute_twiddles(ri, rl)
    _, _, layers = orc.lde_commit(trace, blow, otw, oitw)
    want = layers[0][0].tobytes()
stream = torch.cuda.Stream()
with torch.cuda.stream(stream):
    for ownership in ("contiguous", "halves"):
        mine = np.concatenate([trace[a:a + c] for a, c in owned_columns(rank, world, n_cols, ownership)])
        shard = torch.from_numpy(mine.view(np.int32)).cuda().view(-1)
        sc = PeerShardedCommit(local, stream, n_cols, log, blow, overlap=(groups, peer_sms), ownership=ownership)
        for rep in range(4):                                   # repeated steps: the flags count steps, buffers are reused
            r = sc.step(shard, read_root=(rep % 2 == 0))
            if rank == 0:
                got = r if r is not None else sc.root_bytes()
                assert got == want, (ownership, "step", rep, got.hex(), want.hex())
        ev = sc.step_phases(shard)
        torch.cuda.synchronize()
        assert all(ev[i].elapsed_time(ev[i + 1]) >= 0 for i in range(4))
        if rank == 0:
            assert sc.root_bytes() == want, (ownership, "step_phases")
        host = torch.from_numpy(mine.view(np.int32)).pin_memory()
        stage = torch.empty((n_cols // world, 1 << log), dtype=torch.int32, device="cuda")
        for chunks in (2, 4):
            r = sc.step_from_host(host, stage, chunks)
            if rank == 0:
                assert r == want, (ownership, "step_from_host", chunks)
            r = sc.step_pipelined(shard, chunks)
            if rank == 0:
                assert r == want, (ownership, "step_pipelined", chunks)
        r = sc.step(shard)                                     # a split step right after unsplit ones
        if rank == 0:
            assert r == want, (ownership, "step after chunked steps")
        sc.close()
torch.cuda.synchronize()
dist.barrier()
if rank == 0:
    print("PEER COMMIT OK", world, want.hex())
dist.destroy_process_group()

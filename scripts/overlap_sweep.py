"""Sweep of (column groups, SMs of the NVLink-bound pass) for PeerShardedCommit's overlapped exchange; run under torchrun.
Every configuration must produce the same root; times are CUDA events on the launching stream, max over ranks.
usage: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/overlap_sweep.py [g,k g,k ...]"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from tstwo_b200 import _lib
from tstwo_b200.distributed import PeerShardedCommit

P = 2147483647
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
L = _lib.init(local)
stream = torch.cuda.Stream(device=local); torch.cuda.set_stream(stream)
_lib.check(L.tstwo_set_stream(C.c_void_p(stream.cuda_stream)))
n_cols, log, blow = 256, 22, 1
lc = n_cols // world
g = torch.Generator(device="cuda"); g.manual_seed(1234 + rank)
trace = torch.randint(0, P, (lc, 1 << log), dtype=torch.int32, device="cuda", generator=g).view(-1)
sc = PeerShardedCommit(local, stream, n_cols, log, blow)
cfgs = [tuple(int(x) for x in a.split(",")) for a in sys.argv[1:]] or [(1, 0), (2, 48), (4, 48), (4, 64), (8, 64)]
roots = {}
for cfg in cfgs:                                   # "groups,peer_sms[,taper percent]"
    sc.overlap = None if cfg[0] <= 1 else cfg[:2]
    _lib.check(L.tstwo_debug_set_overlap_taper(cfg[2] if len(cfg) > 2 else 100))
    for _ in range(3):
        sc.step(trace, read_root=False)
    torch.cuda.synchronize(); dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    K = 10
    torch.cuda.synchronize(); dist.barrier()
    a.record(stream)
    for _ in range(K):
        sc.step(trace, read_root=False)          # steps are enqueued back to back; the ranks order themselves with peer flags
    b.record(stream)
    torch.cuda.synchronize(); dist.barrier()
    r = sc.root_bytes()
    t = torch.tensor([a.elapsed_time(b) / K], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        roots[cfg] = r.hex()
        print(f"N={world} groups={cfg[0]:2d} peer_sms={cfg[1]:3d} taper={cfg[2] if len(cfg) > 2 else 100:3d}  {t.item():7.3f} ms/step  root {r.hex()[:16]}", flush=True)
if rank == 0:
    assert len(set(roots.values())) == 1, f"roots differ: {roots}"
    print("all roots equal")
sc.close()
dist.destroy_process_group()

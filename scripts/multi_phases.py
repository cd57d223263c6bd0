"""Per-phase CUDA-event timing of the multi-GPU step (PeerShardedCommit) on every rank; run under torchrun.
usage: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/multi_phases.py"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from tstwo_b200 import _lib
from tstwo_b200.distributed import PeerShardedCommit, top_layers

P = 2147483647
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
L = _lib.init(local)
stream = torch.cuda.Stream(device=local); torch.cuda.set_stream(stream)
_lib.check(L.tstwo_set_stream(C.c_void_p(stream.cuda_stream)))
n_cols, log, blow = 256, 22, 1
lc = n_cols // world
trace = torch.randint(0, P, (lc, 1 << log), dtype=torch.int32, device="cuda")
sc = PeerShardedCommit(local, stream, n_cols, log, blow)
ev = lambda: torch.cuda.Event(enable_timing=True)
GROUPS, PEER_SMS = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (4, 62)

def step(rec=None):
    chk = _lib.check
    e = [ev() for _ in range(5)]
    e[0].record(stream)
    chk(sc.L.tstwo_lde_columns_to_peers_overlapped(trace.data_ptr(), sc.lc, sc.log_size, sc.blow, sc.tw, 0, sc.work.data_ptr(), sc._bases_arr,
                                                   sc.world, sc.rank * sc.lc, 0, GROUPS, PEER_SMS))
    e[1].record(stream)
    dist.all_reduce(sc.flag)
    e[2].record(stream)
    rows = 1 << sc.sub
    ptrs = _lib.dptr_array([sc.recv_ptr + 4 * rows * c for c in range(sc.n_cols)])
    logs = (C.c_int * sc.n_cols)(*([sc.sub] * sc.n_cols))
    chk(sc.L.tstwo_merkle_commit(ptrs, logs, sc.n_cols, sc.layers.data_ptr(), None))
    e[3].record(stream)
    dist.all_gather_into_tensor(sc.roots, sc.layers[:8].contiguous())
    if sc.rank == 0:
        top = top_layers(sc.ops, sc.roots, sc.world)
        sc.ops.to_host_bytes(top[-1] if top else sc.roots)
    e[4].record(stream)
    if rec is not None:
        rec.append(e)

for _ in range(3):
    step()
torch.cuda.synchronize(); dist.barrier()
rec = []
for _ in range(6):
    step(rec)
torch.cuda.synchronize()
names = ["lde + peer stores", "all_reduce (stores visible)", "merkle subtree", "gather roots + top layers"]
tot = [sum(r[i].elapsed_time(r[i + 1]) for r in rec) / len(rec) for i in range(4)]
whole = rec[0][0].elapsed_time(rec[-1][4]) / len(rec)
t = torch.tensor(tot + [whole], device="cuda")
mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
mn = t.clone(); dist.all_reduce(mn, op=dist.ReduceOp.MIN)
if rank == 0:
    print(f"N={world} groups={GROUPS} peer_sms={PEER_SMS}")
    for i, nm in enumerate(names + ["whole step"]):
        print(f"{nm:32s} min over ranks {mn[i].item():7.3f} ms   max {mx[i].item():7.3f} ms")
sc.close()
dist.destroy_process_group()

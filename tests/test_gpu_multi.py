"""The multi-GPU fan-out INSIDE the library, driven by one host thread of one process (SURVEY 8b "Conventions"; the reference
host is one single-threaded process: vcs/prover.ts:13-30 is called inline from prover/index.ts:582-662):
tstwo_init_devices + tstwo_lde_commit_multi (host trace) / tstwo_lde_commit_multi_dev (device shards).  The root must equal
the ORACLE's root of the same trace, and tstwo_lde_commit's on one device, for every device count the box offers
(G = 1 runs everywhere; 2 / 4 / 8 are skipped when the box has fewer GPUs).  Each case runs in a child process because
tstwo_init_devices re-initialises the library."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r'''
import ctypes as C, sys, os
sys.path.insert(0, %(root)r)
import numpy as np, torch
import oracle as orc
from tstwo_b200 import _lib
G, n_cols, log, blow = %(G)d, %(n_cols)d, %(log)d, %(blow)d
L = _lib.load(); chk = _lib.check
ids = (C.c_int * G)(*range(G))
chk(L.tstwo_init_devices(G, ids))
assert L.tstwo_device_count() == G
P = 2**31 - 1
rng = np.random.default_rng(1234 + G)
trace = rng.integers(0, P, (n_cols, 1 << log), dtype=np.uint32)
ri, rl = orc.canonic_half_coset(log + blow)
otw, oitw = orc.precompute_twiddles(ri, rl)
_, _, layers = orc.lde_commit(trace, blow, otw, oitw)
want = layers[0][0].tobytes()
# (a) host trace (pinned), three steps: the events that fence one step against the next are exercised
host = torch.from_numpy(trace.view(np.int32)).pin_memory()
root = (C.c_uint8 * 32)()
for rep in range(3):
    chk(L.tstwo_lde_commit_multi(C.c_void_p(host.data_ptr()), n_cols, log, blow, root, 0))
    assert bytes(root) == want, ("host-fed multi root differs from the oracle", rep)
# pageable host memory works too
chk(L.tstwo_lde_commit_multi(trace.ctypes.data_as(C.c_void_p), n_cols, log, blow, root, 0))
assert bytes(root) == want
# (b) device-resident shards, one tensor per device
lc = n_cols // G
shards = [torch.from_numpy(trace[d * lc:(d + 1) * lc].view(np.int32)).to(f"cuda:{d}") for d in range(G)]
for d in range(G):
    torch.cuda.synchronize(d)
ptrs = (C.c_uint64 * G)(*[s.data_ptr() for s in shards])
for rep in range(3):
    root2 = (C.c_uint8 * 32)()
    chk(L.tstwo_lde_commit_multi_dev(ptrs, n_cols, log, blow, root2, 0))
    assert bytes(root2) == want, ("device-shard multi root differs from the oracle", rep)
# (c) the single-device entry points still work on every slot after tstwo_set_device, and agree
for slot in range(G):
    chk(L.tstwo_set_device(slot))
    tw = C.c_uint64()
    chk(L.tstwo_twiddles_precompute(ri, rl, C.byref(tw)))
    n, ne = 1 << log, 1 << (log + blow)
    dev = f"cuda:{slot}"
    t = torch.from_numpy(trace.view(np.int32)).to(dev)
    ev = torch.empty((n_cols, ne), dtype=torch.int32, device=dev)
    ly = torch.empty(((2 * ne) - 1) * 8, dtype=torch.int32, device=dev)
    torch.cuda.synchronize(slot)
    r1 = (C.c_uint8 * 32)()
    chk(L.tstwo_lde_commit(t.data_ptr(), n_cols, log, blow, tw.value, 0, ev.data_ptr(), ly.data_ptr(), r1, 0))
    assert bytes(r1) == want, ("single-device root on slot", slot)
    # a tree built on one slot is rejected on another
    if G > 1:
        chk(L.tstwo_set_device((slot + 1) %% G))
        rc = L.tstwo_lde_commit(t.data_ptr(), n_cols, log, blow, tw.value, 0, ev.data_ptr(), ly.data_ptr(), r1, 0)
        assert rc != 0 and b"twiddle" in L.tstwo_last_error()
        chk(L.tstwo_set_device(slot))
    chk(L.tstwo_twiddles_free(tw.value))
# bad shapes are refused
assert L.tstwo_lde_commit_multi(C.c_void_p(host.data_ptr()), n_cols + 1 if G > 1 else 0, log, blow, root, 0) != 0
chk(L.tstwo_shutdown())
print("MULTI OK", G, want.hex())
'''


# n_cols a multiple of 32: the host-fed call takes the column-interleaved ownership (first half of every leaf hashed while the
# second half of the trace is still being copied); the other shapes keep contiguous column ranges
@pytest.mark.parametrize("G,n_cols,log,blow", [(1, 6, 15, 1), (1, 32, 15, 1), (2, 12, 16, 1), (2, 64, 16, 1), (4, 16, 17, 1), (4, 64, 16, 1),
                                               (8, 32, 17, 2)])
def test_lde_commit_multi_in_one_process(G, n_cols, log, blow):
    import torch
    if torch.cuda.device_count() < G:
        pytest.skip(f"needs {G} GPUs")
    code = CHILD % {"root": ROOT, "G": G, "n_cols": n_cols, "log": log, "blow": blow}
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "MULTI OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]

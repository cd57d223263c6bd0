"""One-shot leaf layer + tree (tstwo_merkle_commit) against the same tree hashed in two column chunks (tstwo_merkle_leaf_absorb x 2 +
tstwo_merkle_commit_nodes) on one device: a rank's share at N = 8 (256 columns x 2^20 rows) and at N = 2 (x 2^22).  CUDA events."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from tstwo_b200 import _lib

L, chk = _lib.init(0), _lib.check
stream = torch.cuda.Stream()
chk(L.tstwo_set_stream(C.c_void_p(stream.cuda_stream)))
n_cols = 256
for log in (int(a) for a in (sys.argv[1:] or ["20", "22"])):
    rows = 1 << log
    with torch.cuda.stream(stream):
        block = torch.randint(0, 2**31 - 1, (n_cols, rows), dtype=torch.int32, device="cuda")
        a = torch.empty(8 * ((2 << log) - 1), dtype=torch.int32, device="cuda")
        b = torch.empty_like(a)
        ptrs = _lib.dptr_array([block.data_ptr() + 4 * rows * c for c in range(n_cols)])
        logs = (C.c_int * n_cols)(*([log] * n_cols))

        def one_shot():
            chk(L.tstwo_merkle_commit(ptrs, logs, n_cols, a.data_ptr(), None))

        def chunked():
            chk(L.tstwo_merkle_leaf_absorb(block.data_ptr(), rows, 0, 128, n_cols, log, b.data_ptr()))
            chk(L.tstwo_merkle_leaf_absorb(block.data_ptr() + 4 * rows * 128, rows, 128, 128, n_cols, log, b.data_ptr()))
            chk(L.tstwo_merkle_commit_nodes(log, b.data_ptr(), None))

        for name, fn in (("one-shot", one_shot), ("two chunks", chunked)):
            for _ in range(3):
                fn()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(10):
                fn()
            e1.record(stream)
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 10
            comp = rows * 16 + rows - 1
            print(f"256 columns x 2^{log} rows  {name:10s}  {ms:7.3f} ms  {comp / ms / 1e6:6.2f} G compressions/s")
        assert torch.equal(a, b), "chunked tree differs"

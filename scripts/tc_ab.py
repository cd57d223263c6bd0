"""A/B of the tensor-core top stage per pass type on the headline job (step time and per-pass CUDA-event times).
usage: python scripts/tc_ab.py   (tags: bit 0 inverse, bit 1 strided, bit 2 tensor-core stage)"""
import ctypes as C, sys, os
sys.path.insert(0, os.getcwd())
import torch, bench
from tstwo_b200 import _lib
L=_lib.init(0); chk=_lib.check
stream=torch.cuda.Stream(); torch.cuda.set_stream(stream); chk(L.tstwo_set_stream(C.c_void_p(stream.cuda_stream)))
n_cols,log,blow=256,22,1; n=1<<log; ne=2<<log
tw=C.c_uint64(); chk(L.tstwo_twiddles_precompute(1<<(31-log-2), log, C.byref(tw)))
dev=torch.device('cuda',0)
trace=bench.make_trace(torch,dev,0,1,n_cols,n)
coeffs=torch.empty((n_cols,n),dtype=torch.int32,device=dev); evals=torch.empty((n_cols,ne),dtype=torch.int32,device=dev)
layers=torch.empty(((2<<(log+blow))-1)*8,dtype=torch.int32,device=dev); root=(C.c_uint8*32)()
def step(): chk(L.tstwo_lde_commit(trace.data_ptr(), n_cols, log, blow, tw.value, coeffs.data_ptr(), evals.data_ptr(), layers.data_ptr(), root, 0))
for mask in (0,1,5,2,7,0,1):
    L.tstwo_debug_set_fft_tc(mask)
    for _ in range(3): step()
    torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(10): step()
    e1.record(stream); torch.cuda.synchronize()
    L.tstwo_debug_pass_timing(1); step(); tags=(C.c_int*16)(); ms=(C.c_float*16)(); c=L.tstwo_debug_pass_times(16,tags,ms); L.tstwo_debug_pass_timing(0)
    print('tc mask',mask,'step ms %.3f'%(e0.elapsed_time(e1)/10), bytes(root).hex()[:8], ' '.join('%d:%.3f'%(tags[i],ms[i]) for i in range(c)))

"""Phase timeline of the persistent FFT passes (debug aid, GPU box): builds a -DTSTWO_FFT_TRACE variant of the library,
runs one interpolate + evaluate and prints, for CTA 0, what each of the two thread groups was doing when (SM clocks).
usage: python scripts/fft_trace.py [build|run] [n_cols] [log_size]"""
import ctypes as C, os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CS = os.path.join(ROOT, "tstwo_b200", "csrc")
SO = os.path.join(ROOT, "tstwo_b200", "libtstwo_cuda_trace.so")

def build():
    subprocess.check_call(["bash", os.path.join(CS, "build.sh")])
    flags = "-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -ccbin /usr/bin/g++".split()
    os.makedirs(os.path.join(CS, "_obj_trace"), exist_ok=True)
    obj = os.path.join(CS, "_obj_trace", "fft_tma.o")
    subprocess.check_call(["nvcc", *flags, "-DTSTWO_FFT_TRACE", "-c", os.path.join(CS, "fft_tma.cu"), "-o", obj])
    others = [os.path.join(CS, "_obj", f) for f in os.listdir(os.path.join(CS, "_obj")) if f.endswith(".o") and f != "fft_tma.o"]
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-ccbin", "/usr/bin/g++", "-o", SO, obj, *others, "-lcudart"])

def run(n_cols, log):
    L = C.CDLL(SO)
    def chk(rc):
        if rc: raise RuntimeError(L.tstwo_last_error())
    L.tstwo_last_error.restype = C.c_char_p
    chk(L.tstwo_init(0))
    n, ne = 1 << log, 2 << log
    def malloc(b):
        p = C.c_uint64(); chk(L.tstwo_malloc(C.c_uint64(b), C.byref(p))); return p.value
    trace, coeffs, evals = malloc(4 * n * n_cols), malloc(4 * n * n_cols), malloc(4 * ne * n_cols)
    chk(L.tstwo_memset_zero(C.c_uint64(trace), C.c_uint64(4 * n * n_cols)))
    tw = C.c_uint64()
    chk(L.tstwo_twiddles_precompute(C.c_uint32(1 << (31 - log - 2)), log, C.byref(tw)))
    TB = 4096 + 2 * 160
    tb = malloc(TB * 8)
    names = ["wait", "got", "stages", "sync+drain+refill"]
    root = (C.c_uint8 * 32)()
    for which in range(4):
        # passes run in the order inv-contig, inv-strided, fwd-strided, fwd-contig: trace one at a time by arming the
        # buffer, running everything, and keeping only the launch we want (the buffer is overwritten by every launch,
        # so re-run with the trace armed and read after the wanted call)
        pass
    for label, call in (("interpolate", 0), ("evaluate", 1)):
        chk(L.tstwo_memset_zero(C.c_uint64(tb), C.c_uint64(TB * 8)))
        cols_t = (C.c_uint64 * n_cols)(*[trace + 4 * n * c for c in range(n_cols)])
        cols_c = (C.c_uint64 * n_cols)(*[coeffs + 4 * n * c for c in range(n_cols)])
        cols_e = (C.c_uint64 * n_cols)(*[evals + 4 * ne * c for c in range(n_cols)])
        chk(L.tstwo_debug_fft_trace(C.c_uint64(tb)))
        chk(L.tstwo_debug_fft_trace_cta(int(os.environ.get("TRACE_CTA", "0"))))
        if call == 0:
            chk(L.tstwo_interpolate_columns(cols_t, cols_c, n_cols, log, C.c_uint32(1 << (31 - (log - 1) - 2)), tw, 0))
        else:
            chk(L.tstwo_evaluate_columns(cols_c, log, cols_e, n_cols, log + 1, C.c_uint32(1 << (31 - log - 2)), tw, 0))
        chk(L.tstwo_sync())
        host = np.zeros(TB, dtype=np.int64)
        chk(L.tstwo_download(host.ctypes.data_as(C.c_void_p), C.c_uint64(tb), C.c_uint64(host.nbytes)))
        t = host[:2048].reshape(2, 64, 16)
        print(f"== {label}: LAST pass launched (CTA 0), clocks relative to group 0 tile 0 start; columns = phase durations")
        t0 = t[0, 0, 0]
        NP = len(names)
        print("grp tile   start " + " ".join(f"{x:>10s}" for x in names[1:]) + "    total")
        for g in range(2):
            st = host[2048 + 1024 * g: 2048 + 1024 * (g + 1)]
            st = st[st != 0]
            if len(st) > 2:
                d = np.diff(st)
                print(f"group {g}: {len(st)} tiles, clocks per tile (mean over blocks of 32):", [int(d[a:a + 32].mean()) for a in range(0, len(d), 32)])
        ct = host[4096:4096 + 2 * 148].reshape(148, 2)
        dur = (ct[:, 1] - ct[:, 0]) / 1e3
        beg = (ct[:, 0] - ct[:, 0].min()) / 1e3
        print(f"per-CTA duration us: min {dur.min():.0f} max {dur.max():.0f} mean {dur.mean():.0f}; start skew max {beg.max():.0f} us; slowest CTAs {np.argsort(-dur)[:8].tolist()}")
        for k in range(12):
            for g in range(2):
                r = t[g, k]
                if r[NP - 1] == 0: continue
                d = [r[j] - r[j - 1] for j in range(1, NP)]
                print(f"{g:3d} {k:4d} {r[0] - t0:8d} " + " ".join(f"{x:10d}" for x in d) + f" {r[NP - 1] - r[0]:8d}")

if __name__ == "__main__":
    mode = sys.argv[1] if len(sys.argv) > 1 else "run"
    if mode == "build": build()
    else: run(int(sys.argv[2]) if len(sys.argv) > 2 else 256, int(sys.argv[3]) if len(sys.argv) > 3 else 22)

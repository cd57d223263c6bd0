"""ctypes binding of libtstwo_cuda.so (include/tstwo_cuda.h).  No CPU fallback: importing works without a GPU
(so the symbol table can be checked), but any compute call raises unless a B200 is present."""
from __future__ import annotations

import ctypes as C
import os

_DIR = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_DIR, "libtstwo_cuda.so")

dptr = C.c_uint64
u32p = C.POINTER(C.c_uint32)
dptrp = C.POINTER(C.c_uint64)
intp = C.POINTER(C.c_int)

FLAG_QUIRKS = 1

ERR_NAMES = {-1: "BAD_ARG", -2: "SIZE_MISMATCH", -3: "TWIDDLE_MISMATCH", -4: "CUDA", -5: "OOM", -6: "NOT_POW2", -7: "NOT_INIT"}


class TstwoError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"[{ERR_NAMES.get(code, code)}] {msg}")
        self.code = code
        self.msg = msg


class TwiddleMismatch(TstwoError):
    pass


# name -> (restype, argtypes); mirrors include/tstwo_cuda.h one to one
SIGNATURES = {
    "tstwo_init": (C.c_int, [C.c_int]),
    "tstwo_init_devices": (C.c_int, [C.c_int, intp]),
    "tstwo_device_count": (C.c_int, []),
    "tstwo_set_device": (C.c_int, [C.c_int]),
    "tstwo_lde_commit_multi": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_uint8), C.c_uint32]),
    "tstwo_lde_commit_multi_dev": (C.c_int, [dptrp, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_uint8), C.c_uint32]),
    "tstwo_shutdown": (C.c_int, []),
    "tstwo_last_error": (C.c_char_p, []),
    "tstwo_set_stream": (C.c_int, [C.c_void_p]),
    "tstwo_sync": (C.c_int, []),
    "tstwo_mem_trim": (C.c_int, []),
    "tstwo_device_info": (C.c_int, [intp, intp, intp, C.POINTER(C.c_uint64)]),
    "tstwo_launch_count": (C.c_uint64, []),
    "tstwo_malloc": (C.c_int, [C.c_uint64, dptrp]),
    "tstwo_free": (C.c_int, [dptr]),
    "tstwo_memset_zero": (C.c_int, [dptr, C.c_uint64]),
    "tstwo_upload": (C.c_int, [dptr, C.c_void_p, C.c_uint64]),
    "tstwo_download": (C.c_int, [C.c_void_p, dptr, C.c_uint64]),
    "tstwo_grind": (C.c_int, [C.POINTER(C.c_uint8), C.c_uint32, C.POINTER(C.c_uint64)]),
    "tstwo_gather_words": (C.c_int, [C.POINTER(dptr), C.c_int, C.c_int, C.POINTER(C.c_uint32)]),
    "tstwo_copy": (C.c_int, [dptr, dptr, C.c_uint64]),
    "tstwo_bit_reverse": (C.c_int, [dptr, C.c_uint64]),
    "tstwo_twiddles_precompute": (C.c_int, [C.c_uint32, C.c_int, dptrp]),
    "tstwo_twiddles_free": (C.c_int, [C.c_uint64]),
    "tstwo_twiddles_info": (C.c_int, [C.c_uint64, u32p, intp, dptrp, dptrp]),
    "tstwo_interpolate_columns": (C.c_int, [dptrp, dptrp, C.c_int, C.c_int, C.c_uint32, C.c_uint64, C.c_uint32]),
    "tstwo_evaluate_columns": (C.c_int, [dptrp, C.c_int, dptrp, C.c_int, C.c_int, C.c_uint32, C.c_uint64, C.c_uint32]),
    "tstwo_eval_at_point": (C.c_int, [dptr, C.c_int, u32p, u32p, u32p]),
    "tstwo_merkle_commit_layer": (C.c_int, [C.c_int, dptr, dptrp, C.c_int, dptr]),
    "tstwo_merkle_commit": (C.c_int, [dptrp, intp, C.c_int, dptr, C.POINTER(C.c_uint8)]),
    "tstwo_merkle_leaf_absorb": (C.c_int, [dptr, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_int, dptr]),
    "tstwo_merkle_commit_nodes": (C.c_int, [C.c_int, dptr, C.POINTER(C.c_uint8)]),
    "tstwo_fold_line": (C.c_int, [dptrp, C.c_int, C.c_uint32, u32p, dptrp, C.c_uint64]),
    "tstwo_fold_circle_into_line": (C.c_int, [dptrp, dptrp, C.c_int, C.c_uint32, u32p, C.c_uint64]),
    "tstwo_fri_decompose": (C.c_int, [dptrp, C.c_int, dptrp, u32p]),
    "tstwo_fri_commit_layers": (C.c_int, [dptrp, C.c_int, C.c_uint32, C.c_int, dptrp, intp, u32p, C.c_int, C.c_uint64, u32p, dptrp, dptrp,
                                           C.POINTER(C.c_uint8), u32p, intp]),
    "tstwo_accumulate": (C.c_int, [dptrp, dptrp, C.c_uint64]),
    "tstwo_accumulate_quotients": (C.c_int, [C.c_int, C.c_uint32, dptrp, C.c_int, u32p, C.c_int, u32p, intp, intp, u32p, dptrp]),
    "tstwo_scatter_shards": (C.c_int, [C.c_uint64, C.c_int, C.c_uint64, dptrp, C.c_int, C.c_int]),
    "tstwo_scatter_join": (C.c_int, []),
    "tstwo_eval_at_point_many": (C.c_int, [dptrp, C.c_int, C.c_int, u32p, u32p, u32p]),
    "tstwo_eval_at_point_noalloc": (C.c_int, [dptr, C.c_int, u32p, u32p, u32p]),
    "tstwo_eval_at_point_noalloc_max_log": (C.c_int, []),
    "tstwo_merkle_decommit": (C.c_int, [C.c_uint64, C.c_int, dptrp, intp, C.c_int, intp, u32p, u32p, C.POINTER(C.c_uint8), u32p, C.POINTER(C.c_uint64)]),
    "tstwo_fri_decommit_layers": (C.c_int, [C.c_int, dptrp, dptrp, intp, u32p, C.c_int, u32p, C.c_uint64, C.POINTER(C.c_uint8), C.c_uint64,
                                            u32p, C.c_uint64, C.POINTER(C.c_uint64)]),
    "tstwo_widefib_constraints": (C.c_int, [C.c_int, C.c_int, dptrp, C.c_int, u32p, u32p, dptrp]),
    "tstwo_accumulate_quotients_ex": (C.c_int, [C.c_int, C.c_uint32, dptrp, C.c_int, u32p, C.c_int, u32p, intp, intp, u32p, dptrp, C.c_uint32]),
    "tstwo_lde_commit": (C.c_int, [dptr, C.c_int, C.c_int, C.c_int, C.c_uint64, dptr, dptr, dptr, C.POINTER(C.c_uint8), C.c_uint32]),
    "tstwo_lde_columns_sharded": (C.c_int, [dptr, C.c_int, C.c_int, C.c_int, C.c_uint64, dptr, dptr, C.c_int, C.c_uint32]),
    "tstwo_lde_columns_to_peers": (C.c_int, [dptr, C.c_int, C.c_int, C.c_int, C.c_uint64, dptr, dptr, dptrp, C.c_int, C.c_int, C.c_uint32]),
    "tstwo_lde_columns_to_peers_overlapped": (C.c_int, [dptr, C.c_int, C.c_int, C.c_int, C.c_uint64, dptr, dptr, dptrp, C.c_int, C.c_int,
                                                         C.c_uint32, C.c_int, C.c_int]),
    "tstwo_lde_columns_to_peers_halves": (C.c_int, [dptr, C.c_int, C.c_int, C.c_int, C.c_uint64, dptr, dptr, dptrp, C.c_int, C.c_int, C.c_int,
                                                     C.c_uint32, C.c_int, C.c_int, dptrp, dptrp, C.c_int, C.c_uint32]),
    "tstwo_peer_join": (C.c_int, []),
    "tstwo_ipc_export": (C.c_int, [dptr, C.POINTER(C.c_uint8)]),
    "tstwo_ipc_import": (C.c_int, [C.POINTER(C.c_uint8), C.POINTER(C.c_uint64)]),
    "tstwo_ipc_close": (C.c_int, [dptr]),
    "tstwo_peer_signal": (C.c_int, [C.POINTER(C.c_uint64), C.c_int, C.c_uint32]),
    "tstwo_peer_wait": (C.c_int, [dptr, C.c_int, C.c_uint32, C.c_uint32]),
    "tstwo_peer_put": (C.c_int, [dptr, dptr, C.c_int, dptr, C.c_uint32]),
    "tstwo_lde_commit_host": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_uint64, dptr, dptr, dptr, dptr,
                                        C.POINTER(C.c_uint8), C.c_uint32]),
}

# include/tstwo_cuda_debug.h: test / measurement hooks, not part of the drop-in boundary
DEBUG_SIGNATURES = {
    "tstwo_debug_set_fft_stream": (C.c_int, [C.c_int]),
    "tstwo_debug_force_naive_fft": (C.c_int, [C.c_int]),
    "tstwo_debug_set_fft_rows9": (C.c_int, [C.c_int]),
    "tstwo_debug_set_fft_radix": (C.c_int, [C.c_int]),
    "tstwo_debug_last_fft_mode": (C.c_int, []),
    "tstwo_debug_set_fft_tc": (C.c_int, [C.c_int]),
    "tstwo_debug_set_overlap_taper": (C.c_int, [C.c_int]),
    "tstwo_debug_pass_timing": (C.c_int, [C.c_int]),
    "tstwo_debug_pass_times": (C.c_int, [C.c_int, intp, C.POINTER(C.c_float)]),
    "tstwo_debug_set_leaf_mode": (C.c_int, [C.c_int]),
    "tstwo_debug_last_leaf_mode": (C.c_int, []),
    "tstwo_debug_set_quot_mode": (C.c_int, [C.c_int]),
    "tstwo_debug_last_quot_mode": (C.c_int, []),
    "tstwo_debug_field_op": (C.c_int, [C.c_int, C.c_int, u32p, u32p, u32p, C.c_int]),
}

_lib = None


def load():
    """dlopen the library and bind every symbol of the header.  Raises if the .so is missing (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise RuntimeError(f"{SO_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(tstwo_b200 has no CPU fallback)")
        lib = C.CDLL(SO_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)   # AttributeError if the symbol is missing
            fn.restype = res
            fn.argtypes = args
        for name, (res, args) in DEBUG_SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc: int):
    if rc != 0:
        msg = load().tstwo_last_error().decode(errors="replace")
        if rc == -3:
            raise TwiddleMismatch(rc, msg)
        raise TstwoError(rc, msg)


_inited = None


def init(device: int = 0):
    global _inited
    lib = load()
    if _inited != device:
        check(lib.tstwo_init(device))
        _inited = device
    return lib


def dptr_array(addrs):
    return (C.c_uint64 * max(1, len(addrs)))(*addrs)


def u32_array(vals):
    return (C.c_uint32 * max(1, len(vals)))(*[int(v) for v in vals])

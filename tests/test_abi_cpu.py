"""CPU-side checks: the C-ABI library loads and exports every symbol include/tstwo_cuda.h declares, fails loudly
without a GPU (no CPU fallback), and the host index arithmetic mirrors the reference (vs the oracle)."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_header_symbol():
    import __graft_entry__ as g
    if not os.path.exists(os.path.join(ROOT, "tstwo_b200", "libtstwo_cuda.so")):
        g.build()
    from tstwo_b200 import _lib
    lib = _lib.load()
    hdr = open(os.path.join(ROOT, "include", "tstwo_cuda.h")).read()
    names = set(re.findall(r"\b(tstwo_[a-z0-9_]+)\s*\(", hdr))
    assert len(names) >= 28
    for n in names:
        assert hasattr(lib, n), n
        assert n in _lib.SIGNATURES, n
    # the test / measurement hooks are declared too (include/tstwo_cuda_debug.h), and nothing named tstwo_* is exported
    # without a declaration in one of the two headers
    dbg = open(os.path.join(ROOT, "include", "tstwo_cuda_debug.h")).read()
    dnames = set(re.findall(r"\b(tstwo_debug_[a-z0-9_]+)\s*\(", dbg))
    assert dnames == set(_lib.DEBUG_SIGNATURES), dnames ^ set(_lib.DEBUG_SIGNATURES)
    for n in dnames:
        assert hasattr(lib, n), n
    import subprocess
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.SO_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (tstwo_[a-z0-9_]+)$", out, flags=re.M))
    assert exported, "nm found no exports"
    undeclared = exported - names - dnames
    assert not undeclared, f"exported but declared in no header: {sorted(undeclared)}"


def _header_decls():
    """{name: (number of parameters, FFI return type)} of every function include/tstwo_cuda.h declares."""
    h = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "tstwo_cuda.h")).read(), flags=re.S)
    out = {}
    for ret, name, args in re.findall(r"\b(int|uint64_t|const char \*)\s*(tstwo_[a-z0-9_]+)\s*\(([^;]*?)\)\s*;", h, flags=re.S):
        params = [a for a in (x.strip() for x in args.split(",")) if a and a != "void"]
        out[name] = (len(params), {"int": "T.i32", "uint64_t": "T.u64", "const char *": "T.cstring"}[ret])
    return out


def test_ts_binding_matches_header():
    """ts/cuda_backend.ts (the bun:ffi binding a TStwo maintainer adds; not executable here, no Bun) declares EVERY entry point
    of the header with the right number of arguments and the right return type -- and nothing the header does not have."""
    decls = _header_decls()
    ts = open(os.path.join(ROOT, "ts", "cuda_backend.ts")).read()
    bound = {m.group(1): (len([x for x in m.group(2).split(",") if x.strip()]), m.group(3))
             for m in re.finditer(r"^\s*(tstwo_[a-z0-9_]+): \{ args: \[([^\]]*)\], returns: (T\.[a-z0-9]+) \}", ts, flags=re.M)}
    assert set(bound) == set(decls), (sorted(set(decls) - set(bound)), sorted(set(bound) - set(decls)))
    for name, (n, ret) in decls.items():
        assert bound[name] == (n, ret), (name, bound[name], (n, ret))
    # and the ctypes table has the same arities
    from tstwo_b200 import _lib
    for name, (n, _) in decls.items():
        assert len(_lib.SIGNATURES[name][1]) == n, name


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from tstwo_b200 import _lib, CudaColumn
    with pytest.raises(_lib.TstwoError, match="no CUDA device|CUDA"):
        CudaColumn.zeros(4)


def test_product_does_not_import_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "tstwo_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "tstwo_oracle" not in src, f


def test_host_circle_mirror(orc):
    from tstwo_b200 import CanonicCoset, CirclePointIndex, Coset
    for log in range(1, 12):
        cc = CanonicCoset.new(log)
        hc = cc.half_coset()
        assert (hc.initial_index.value, hc.log_size) == orc.canonic_half_coset(log)
        dom = cc.circleDomain()
        assert dom.isCanonic()
        for i in [0, 1, dom.size() // 2, dom.size() - 1]:
            assert dom.at(i) == orc.domain_point(hc.initial_index.value, hc.log_size, i)
    big = Coset.half_odds(10)
    assert Coset.half_odds(6).is_doubling_of(big) and not Coset.half_odds(6).shift(CirclePointIndex(2)).is_doubling_of(big)
    assert not big.is_doubling_of(Coset.half_odds(6))
    assert CirclePointIndex(1).to_point() == (2, 1268011823)
    with pytest.raises(ValueError):
        CanonicCoset.new(0)


def test_blake2s_kernels_keep_the_uniform_register_imad_form():
    """The two BLAKE2s kernels that carry the Merkle time (leaf layer, node layers) rely on ptxas emitting their additions as
    IMAD R, R, UR, R (`one` in a uniform register: 1345 clk per warp-compression against 1556 for the three-register form,
    profiles/r01_microbench_blake2s_forms.txt).  The choice is a ptxas heuristic (launch bounds, loop shape, even blockDim.x vs a
    literal flip it), so the built library is disassembled here: a toolchain change that loses the form fails loudly instead of
    costing 14 % silently."""
    import re
    import shutil
    import subprocess
    from tstwo_b200 import _lib
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    sass = subprocess.run(["cuobjdump", "-sass", _lib.SO_PATH], capture_output=True, text=True).stdout
    for kernel in ("merkle_leaf_tma_kernel", "merkle_leaf_chunk_kernel", "merkle_node_kernel", "merkle_node_cols_kernel"):
        body, on = [], False
        for line in sass.splitlines():
            if "Function :" in line:
                on = kernel in line
            elif on:
                body.append(line)
        forms = [m.group(1) for l in body for m in [re.search(r"\bIMAD R\d+, R\d+, (UR\d+|R\d+)[.\w]*, R\d+", l)] if m]
        ur = sum(f.startswith("UR") for f in forms)
        assert len(forms) >= 400, f"{kernel}: expected ~480 IMAD additions per compression, found {len(forms)}"
        assert ur >= 0.95 * len(forms), f"{kernel}: {ur} of {len(forms)} IMAD additions use the uniform-register form"

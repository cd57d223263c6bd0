"""tstwo_eval_at_point_noalloc (no counterpart in the reference; added for callers that cannot
call cudaMalloc): through the C ABI, against the existing tstwo_eval_at_point on the same inputs -- bit-exact, as every integer
result in this repo -- on two calls in a row (the fixed scratch array keeps the first call's contents), at several sizes up to
the stated maximum, and the rejection beyond it.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
P = 2147483647
PX = [1, 0, 478637715, 513582971]
PY = [992285211, 649143431, 740191619, 1186584352]


def _eval_both(tb, L, lib, coeffs, log, px, py):
    col = tb.CudaColumn.from_cpu(coeffs)
    a, b = (C.c_uint32 * 4)(), (C.c_uint32 * 4)()
    lib.check(L.tstwo_eval_at_point_noalloc(col.ptr, log, lib.u32_array(px), lib.u32_array(py), a))
    lib.check(L.tstwo_eval_at_point(col.ptr, log, lib.u32_array(px), lib.u32_array(py), b))
    return list(a), list(b)


def test_noalloc_eval_at_point_matches_existing_entry_point():
    import tstwo_b200 as tb
    from tstwo_b200 import _lib as lib
    L = lib.init(0)
    max_log = L.tstwo_eval_at_point_noalloc_max_log()
    assert max_log == 20
    rng = np.random.default_rng(0x20A110C)
    # two calls in a row at the maximum size with different inputs: the second must not see the first one's partial sums
    for rep in range(2):
        coeffs = rng.integers(0, P, 1 << max_log, dtype=np.uint32)
        px = [int(v) for v in rng.integers(0, P, 4)] if rep else PX
        py = [int(v) for v in rng.integers(0, P, 4)] if rep else PY
        got, want = _eval_both(tb, L, lib, coeffs, max_log, px, py)
        assert got == want, f"call {rep}: {got} != {want}"
    # a large call followed by small ones (stale partials of the large call stay in the array), every level structure
    for log in (0, 1, 2, 3, 4, 5, 7, 8, 9, 12, 13, 16, 17, 19, 20, 6):
        coeffs = rng.integers(0, P, 1 << log, dtype=np.uint32)
        got, want = _eval_both(tb, L, lib, coeffs, log, PX, PY)
        assert got == want, f"log {log}: {got} != {want}"
    # the same call twice gives the same answer
    coeffs = rng.integers(0, P, 1 << 18, dtype=np.uint32)
    first, want = _eval_both(tb, L, lib, coeffs, 18, PX, PY)
    second, _ = _eval_both(tb, L, lib, coeffs, 18, PX, PY)
    assert first == second == want


def test_noalloc_rejects_sizes_beyond_the_fixed_scratch():
    import tstwo_b200 as tb
    from tstwo_b200 import _lib as lib
    L = lib.init(0)
    col = tb.CudaColumn.zeros(1 << 21)
    out = (C.c_uint32 * 4)()
    rc = L.tstwo_eval_at_point_noalloc(col.ptr, 21, lib.u32_array(PX), lib.u32_array(PY), out)
    assert rc != 0
    msg = L.tstwo_last_error()
    assert b"exceeds the fixed scratch" in msg and b"2^20" in msg
    assert L.tstwo_eval_at_point_noalloc(0, 4, lib.u32_array(PX), lib.u32_array(PY), out) != 0
    # still usable after a rejected call
    lib.check(L.tstwo_eval_at_point_noalloc(col.ptr, 20, lib.u32_array(PX), lib.u32_array(PY), out))
    assert list(out) == [0, 0, 0, 0]

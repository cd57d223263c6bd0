"""Environment knobs of the library: TSTWO_TRACE span lines, TSTWO_CUDA_DEVICES as the device list of
tstwo_init_devices(n, NULL), TSTWO_QUIRKS as the host mirror's default.  Child processes: the library reads them once."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r"""
import sys, numpy as np
sys.path.insert(0, %r)
import tstwo_b200 as tb
from tstwo_b200 import _lib
L = _lib.load()
_lib.check(L.tstwo_init_devices(0, None))            # the list comes from TSTWO_CUDA_DEVICES
assert L.tstwo_device_count() == 1
print("quirks", tb.CudaBackend.quirks)
log, blow, n_cols = 14, 1, 4
trace = np.random.default_rng(1).integers(0, 2147483647, (n_cols, 1 << log), dtype=np.uint32)
tw = tb.CudaBackend.precomputeTwiddles(tb.CanonicCoset.new(log + blow).half_coset())
tree = tb.CommitmentTreeProver.from_trace(tb.CudaColumn.from_cpu(trace.reshape(-1)), n_cols, log, blow, tw)
print("root", tree.root().hex())
"""


@pytest.mark.gpu
def test_trace_devices_and_quirks_from_the_environment():
    env = dict(os.environ, TSTWO_TRACE="2", TSTWO_CUDA_DEVICES="0", TSTWO_QUIRKS="0")
    r = subprocess.run([sys.executable, "-c", CHILD % ROOT], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "quirks False" in r.stdout and "root " in r.stdout
    lines = [l for l in r.stderr.splitlines() if l.startswith("tstwo: ")]
    names = " | ".join(lines)
    for span in ("Interpolation for commitment", "Extension", "Merkle"):
        assert span in names, names
    assert all("ms" in l and "(stream drained)" in l for l in lines)
    # without the variable nothing is printed and the quirk default is the reference's behaviour
    env = {k: v for k, v in os.environ.items() if not k.startswith("TSTWO_")}
    env["TSTWO_CUDA_DEVICES"] = "0"
    r = subprocess.run([sys.executable, "-c", CHILD % ROOT], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "quirks True" in r.stdout and "tstwo: " not in r.stderr


def test_device_list_errors_need_no_gpu():
    from tstwo_b200 import _lib
    L = _lib.load()
    old = os.environ.pop("TSTWO_CUDA_DEVICES", None)
    try:
        assert L.tstwo_init_devices(0, None) != 0 and b"TSTWO_CUDA_DEVICES is not set" in L.tstwo_last_error()
        os.environ["TSTWO_CUDA_DEVICES"] = "0,x"
        assert L.tstwo_init_devices(0, None) != 0 and b"cannot parse" in L.tstwo_last_error()
        os.environ["TSTWO_CUDA_DEVICES"] = "0,1"
        assert L.tstwo_init_devices(3, None) != 0 and b"names 2 devices" in L.tstwo_last_error()
    finally:
        os.environ.pop("TSTWO_CUDA_DEVICES", None)
        if old is not None:
            os.environ["TSTWO_CUDA_DEVICES"] = old

"""PeerShardedCommit (tstwo_b200/distributed.py: one process per GPU, the exchange fused into the last FFT pass, release/acquire flags
between the ranks) against the oracle's root under torchrun, for contiguous and column-interleaved ("halves") ownership -- the latter
hashes the first half of every leaf (tstwo_merkle_leaf_absorb) beside the last storing pass.  World size 1 runs everywhere (the
"peer" is the rank itself: same calls, same flags); 2 / 4 are skipped when the box has fewer GPUs."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("G,n_cols,log,blow,groups,peer_sms", [(1, 32, 15, 1, 2, 62), (1, 64, 16, 1, 4, 66), (2, 64, 16, 1, 4, 62),
                                                               (4, 128, 16, 1, 2, 62), (2, 64, 15, 2, 8, 40)])
def test_peer_sharded_commit_both_ownerships(G, n_cols, log, blow, groups, peer_sms):
    import torch
    if torch.cuda.device_count() < G:
        pytest.skip(f"needs {G} GPUs")
    port = 29600 + (os.getpid() + 7 * G + log) % 300
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={G}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "peer_commit_child.py"),
           str(n_cols), str(log), str(blow), str(groups), str(peer_sms)]
    env = dict(os.environ, TSTWO_PEER_TIMEOUT_MS="20000")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)
    assert r.returncode == 0 and "PEER COMMIT OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]

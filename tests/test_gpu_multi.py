"""N > 1 on real GPUs: tests/tools/multi_check.py under torchrun (one rank per GPU, NCCL
all-gather of the packed stage positions overlapped with the local-tile force launch).
Skipped on boxes with fewer than two GPUs (the driver's GPU test box has one); the
world_size-2 logic is covered on CPU by tests/test_multirank_gloo.py."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.skipif(_n_gpus() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("p2p,port", [("0", "29533"), ("1", "29534")])
def test_two_ranks_match_oracle(p2p, port):
    """p2p = 0: NCCL all-gather per stage; p2p = 1: peer-pull exchange (the force kernel's
    bulk-async copies read the other rank's tiles from its HBM; no collective per stage)."""
    env = dict(os.environ, NBODY_P2P=p2p)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                        "--nproc-per-node", "2", "--master-addr", "127.0.0.1", "--master-port",
                        port, os.path.join(ROOT, "tests", "tools", "multi_check.py"), "3000"],
                       capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0 and "MULTI-GPU CHECK PASSED" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]

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
@pytest.mark.parametrize("p2p,sym,port", [("0", "0", "29533"), ("1", "0", "29534"), ("0", "auto", "29536")])
def test_two_ranks_match_oracle(p2p, sym, port):
    """p2p = 0: NCCL all-gather per stage; p2p = 1: peer-pull exchange (the force kernel's
    bulk-async copies read the other rank's tiles from its HBM; no collective per stage);
    sym = "0": the one-sided kernel, "auto": the library's choice (the pair-symmetric kernel at this
    size: blocks of one global schedule split over the ranks, sums reduce-scattered); sym = "1" is
    also exercised from tests/test_zzz_gpu_pair_symmetry.py, which sorts last."""
    env = dict(os.environ, NBODY_P2P=p2p, NBODY_PAIR_SYM=sym)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                        "--nproc-per-node", "2", "--master-addr", "127.0.0.1", "--master-port",
                        port, os.path.join(ROOT, "tests", "tools", "multi_check.py"), "3000"],
                       capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0 and "MULTI-GPU CHECK PASSED" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.skipif(_n_gpus() < 2, reason="needs >= 2 GPUs")
def test_two_ranks_deterministic_row_sums():
    """Option "deterministic" on two ranks: the fixed-point limbs are reduce-scattered as 64-bit
    integers; every rank against the oracle, repeated evaluations and the owned-rows upload path
    must give the same BITS (tests/tools/multi_check.py, NBODY_DETERMINISTIC=1)."""
    env = dict(os.environ, NBODY_P2P="0", NBODY_PAIR_SYM="auto", NBODY_DETERMINISTIC="1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                        "--nproc-per-node", "2", "--master-addr", "127.0.0.1", "--master-port",
                        "29537", os.path.join(ROOT, "tests", "tools", "multi_check.py"), "3000"],
                       capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0 and "deterministic: fixed-point row sums" in r.stdout, \
        r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.skipif(_n_gpus() < 2, reason="needs >= 2 GPUs")
def test_cxx_driver_two_gpus_matches_one(tmp_path):
    """The C++ host driver with one rank-thread per GPU (NCCL from threads of one process)."""
    import numpy as np
    from parallel_comp_labs_b200 import lab
    exe = os.path.join(ROOT, "parallel_comp_labs_b200", "bin", "nbody_b200")
    inp = str(tmp_path / "in.txt")
    assert subprocess.run([exe, "--generate", "uniform", "--n", "9000", "--generate-out", inp]).returncode == 0
    outs = []
    for g in (1, 2):
        fin = str(tmp_path / f"f{g}.txt")
        r = subprocess.run([exe, "--input", inp, "--scheme", "rk4", "--tau", "0.1", "--t-final", "0.25",
                            "--gpus", str(g), "--final-state", fin, "--timing-out", str(tmp_path / "t.csv"),
                            "--traj-out", str(tmp_path / f"traj{g}.csv")],
                           capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr + r.stdout
        outs.append(lab.read_data(fin))
    assert np.abs(outs[0] - outs[1]).max() / np.abs(outs[0][:, 1:]).max() < 1e-12
    a = open(tmp_path / "traj1.csv").read().splitlines()
    b = open(tmp_path / "traj2.csv").read().splitlines()
    assert len(a) == len(b) == 1 + 1 + 3 and a[0] == b[0]      # header, t_start row, 3 steps

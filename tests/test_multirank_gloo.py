"""world_size-2 (and 3) gloo test of the N>1 host logic on CPU: the C ABI's row partition
(nbody_partition), equal-block padded all-gather assembly, and rank-0-only output, driven
through torch.distributed exactly as bench.py / the host mirror drive NCCL on GPUs.
The per-rank compute is played by the ORACLE here (test infrastructure standing in for the
device), so what is checked is the decomposition: P ranks reproduce the single-rank result
bit for bit, as the reference's MPI decomposition does (tests/test_oracle_vs_reference.py)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, steps, out_path):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import Oracle
    from parallel_comp_labs_b200 import capi, generators
    o = Oracle()
    bodies = generators.uniform_bodies(n, seed=11)
    begin, end, blk = capi.partition(n, world, rank)
    n_pad = blk * world
    tau, eps, G = 0.1, 1e-6, 6.67e-11
    # packed {x,y,z,m} replicated array + owner-local velocities, as on the device
    pos = np.zeros((n_pad, 4))
    pos[:n, :3] = bodies[:, 1:4]
    pos[:n, 3] = bodies[:, 0]
    vel = bodies[begin:end, 4:7].copy()

    def gather(p):  # equal-count in-place all-gather of this rank's block
        mine = torch.from_numpy(p[rank * blk:(rank + 1) * blk].copy())
        parts = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(parts, mine)
        for r in range(world):
            p[r * blk:(r + 1) * blk] = parts[r].numpy()

    def acc(p):  # rows [begin,end) against all n_pad bodies (zero-mass pads contribute 0)
        aos = np.zeros((n_pad, 7))
        aos[:, 0] = p[:, 3]
        aos[:, 1:4] = p[:, :3]
        return o.rhs(aos, begin, end, eps=eps, G=G)[begin:end, 4:7]

    for _ in range(steps):  # explicit midpoint, the fused stage kernel's arithmetic
        a1 = acc(pos)
        stage = pos.copy()
        stage[begin:end, :3] = pos[begin:end, :3] + vel * (tau / 2)
        v2 = vel + a1 * (tau / 2)
        gather(stage)
        a2 = acc(stage)
        pos[begin:end, :3] = pos[begin:end, :3] + v2 * tau
        vel = vel + a2 * tau
        gather(pos)
    # velocities are owner-local; gather them for the final state (nbody_download_aos)
    vall = np.zeros((n_pad, 3))
    vall[begin:end] = vel
    vfull = np.zeros((n_pad, 4))
    vfull[:, :3] = vall
    gather(vfull)
    if rank == 0:
        final = np.concatenate([pos[:n, 3:4], pos[:n, :3], vfull[:n, :3]], axis=1)
        np.save(out_path, final)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n", [(2, 300), (3, 1000)])
def test_partitioned_rk2_matches_single_rank(tmp_path, world, n):
    from oracle import Oracle
    from parallel_comp_labs_b200 import generators
    out = str(tmp_path / "final.npy")
    mp.spawn(_worker, args=(world, _free_port(), n, 2, out), nprocs=world, join=True)
    got = np.load(out)
    want, _ = Oracle().rk2(generators.uniform_bodies(n, seed=11), 0.1, 0.0, 1.0, max_steps=2,
                           rows=False)
    assert np.array_equal(got, want)


def test_partition_properties():
    from parallel_comp_labs_b200 import capi
    for n, world in ((4, 1), (4, 4), (10000, 8), (20000, 3), (1 << 20, 8), (1000003, 7)):
        covered = 0
        blks = set()
        for r in range(world):
            b, e, blk = capi.partition(n, world, r)
            assert b == min(n, r * blk) and b <= e <= n and e - b <= blk
            covered += e - b
            blks.add(blk)
        assert covered == n and len(blks) == 1
        blk = blks.pop()
        assert blk % 256 == 0 and blk * world >= n and blk * world - n < 256 * world + world
    with pytest.raises(ValueError, match="greater than the problem dimension"):
        capi.partition(4, 5, 0)

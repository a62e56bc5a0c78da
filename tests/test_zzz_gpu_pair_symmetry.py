"""Parity of the pair-symmetric force kernel (option "pair_symmetry"; the default for n >= 2048:
parallel_comp_labs_b200/csrc/force_sym_kernels.cuh) against the oracle and against the one-sided
kernel.  Same tolerances as tests/test_gpu_parity.py.  Its decomposition is also proven on the CPU
by tests/test_host_logic.py::test_pair_symmetric_schedule_covers_every_pair_once."""
import numpy as np
import pytest

from conftest import rel_inf
from parallel_comp_labs_b200 import capi, generators
from test_gpu_parity import check_state

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300, method="thread")]

TOL64_RHS, TOL64_TRAJ, TOL32 = 1e-12, 1e-10, 1e-4


def sym_rhs(bodies, prec=64, eps=1e-6, G=6.67e-11, sym_variant=-1, slices=0, sym=1, unimass=-1):
    with capi.Context(bodies.shape[0], prec) as ctx:
        ctx.set_params(G, eps)
        ctx.set_option("pair_symmetry", sym)
        ctx.set_option("sym_variant", sym_variant)
        ctx.set_option("slices", slices)
        ctx.set_option("unimass", unimass)
        ctx.upload_aos(bodies)
        out = ctx.rhs().astype(np.float64)
        tm = ctx.timing()
        return out, tm


@pytest.mark.parametrize("n", [1, 2, 3, 255, 256, 257, 511, 512, 513, 1000, 1025, 1537, 2049, 4097])
def test_sym_rhs_ragged_sizes_vs_oracle(oracle, n):
    bodies = generators.uniform_bodies(n, seed=100 + n)
    want = oracle.rhs(bodies, threads=4)[:, 4:7]
    got, tm = sym_rhs(bodies)
    assert tm.pair_symmetry == 1
    assert got.shape == (n, 3) and np.all(np.isfinite(got))
    if n == 1:
        assert np.all(got == 0)
    else:
        assert rel_inf(got, want) < TOL64_RHS


def test_sym_rhs_golden_fixtures(golden):
    got, _ = sym_rhs(golden["body4"])                 # equal masses -> uniform-mass kernel
    assert rel_inf(got, golden["ref_rhs_body4"]) < TOL64_RHS
    got, _ = sym_rhs(golden["rand512"])
    assert rel_inf(got, golden["ref_rhs_rand512"]) < TOL64_RHS
    plu = golden["plummer1000_in"]       # has a pair inside the clamp radius and a coincident pair
    for um in (-1, 0):
        got, tm = sym_rhs(plu, eps=1e-3, unimass=um)
        assert np.all(np.isfinite(got))
        assert rel_inf(got, golden["ref_rhs_plummer1000_eps1e-3"]) < TOL64_RHS
        for i in (10, 11, 20, 21):
            assert rel_inf(got[i], golden["ref_rhs_plummer1000_eps1e-3"][i]) < 1e-11


def test_sym_all_variants_slicings_and_masses(oracle):
    bodies = generators.plummer_bodies(3000, seed=5)
    bodies[7, 1:4] = bodies[8, 1:4] + 1e-5            # planted pair inside the clamp radius
    bodies[1500, 1:4] = bodies[20, 1:4]               # coincident pair in DIFFERENT blocks
    want = oracle.rhs(bodies, eps=1e-3, threads=8)[:, 4:7]
    gen = bodies.copy()
    gen[:, 0] *= np.linspace(0.5, 2.0, gen.shape[0])  # general masses
    gen[33, 0] = 0.0                                  # a massless body feels, does not act
    want_gen = oracle.rhs(gen, eps=1e-3, threads=8)[:, 4:7]
    var = 0
    while True:
        try:
            got, tm = sym_rhs(bodies, eps=1e-3, sym_variant=var)
        except capi.NBodyError:
            break
        assert tm.uniform_mass == 1
        assert rel_inf(got, want) < TOL64_RHS, var
        got, tm = sym_rhs(gen, eps=1e-3, sym_variant=var)
        assert tm.uniform_mass == 0
        assert rel_inf(got, want_gen) < TOL64_RHS, var
        var += 1
    assert var >= 9                                   # 5 shuffled-j + 4 re-read-j builds
    for s in (1, 2, 5):
        got, _ = sym_rhs(bodies, eps=1e-3, slices=s)
        assert rel_inf(got, want) < TOL64_RHS, s
    v32 = 0
    while True:
        try:
            got, _ = sym_rhs(bodies, prec=32, eps=1e-3, sym_variant=v32)
        except capi.NBodyError:
            break
        mask = np.ones(3000, bool)
        mask[[7, 8]] = False                          # their separation itself rounds in FP32
        assert np.abs(got[mask] - want[mask]).max() / np.abs(want[mask]).max() < TOL32, v32
        v32 += 1
    assert v32 >= 5


def test_sym_edge_cases_and_counters(oracle):
    b = generators.uniform_bodies(1400, seed=3)
    b[5, 0] = 0.0
    b[40, 1:4] = b[41, 1:4]             # coincident, same block
    b[100, 1:4] = b[901, 1:4] + 3e-7    # inside eps, different blocks
    want = oracle.rhs(b)[:, 4:7]
    got, tm = sym_rhs(b)
    assert np.all(np.isfinite(got)) and rel_inf(got, want) < TOL64_RHS
    for i in (100, 901):
        assert rel_inf(got[i], want[i]) < 1e-11
    # the counters say what was computed: all N^2 interactions delivered, about half evaluated
    assert tm.interactions == 1400 * 1400
    assert 0.5 * 1400 * 1400 <= tm.pair_evaluations < 1400 * 1400
    dflt, tm0 = sym_rhs(b, sym=0)
    assert tm0.pair_symmetry == 0 and tm0.pair_evaluations == tm0.interactions == 1400 * 1400
    assert rel_inf(got, dflt) < 1e-13
    # eps = 0: the self pair is NaN in the reference and in both kernels
    assert np.all(np.isnan(sym_rhs(b, eps=0.0)[0]))
    # everything inside the clamp radius
    got, _ = sym_rhs(b, eps=1e4)
    assert rel_inf(got, oracle.rhs(b, eps=1e4)[:, 4:7]) < TOL64_RHS


def test_sym_large_n_row_subsample_and_momentum(oracle):
    n = 262144
    bodies = generators.plummer_bodies(n, seed=0x5EED0003)
    got, tm = sym_rhs(bodies, eps=1e-3)
    for begin in (0, 131072 - 8, n - 16):
        want = oracle.rhs(bodies, begin, begin + 16, eps=1e-3, threads=8)[begin:begin + 16, 4:7]
        assert np.abs(got[begin:begin + 16] - want).max() / np.abs(want).max() < TOL64_RHS
        per_row = np.abs(got[begin:begin + 16] - want).max(axis=1) / np.abs(want).max(axis=1)
        assert per_row.max() < 1e-10
    mom = (bodies[:, 0:1] * got).sum(axis=0)
    assert np.abs(mom).max() / (bodies[:, 0].sum() * np.abs(got).max()) < 1e-12
    assert tm.pair_evaluations < 0.51 * n * n


@pytest.mark.parametrize("n,graph", [(4, 1), (777, 1), (3000, 0)])
def test_sym_rk4_matches_default_and_oracle(oracle, golden, n, graph):
    bodies = golden["body4"] if n == 4 else generators.uniform_bodies(n, seed=21)
    want, _ = oracle.rk4(bodies, 0.1, 0.0, 1.0, max_steps=3, rows=False, threads=8)
    res = []
    for sym in (0, 1):
        with capi.Context(n, 64) as ctx:
            ctx.set_option("pair_symmetry", sym)
            ctx.set_option("use_graph", graph)
            ctx.upload_aos(bodies)
            ctx.step(capi.NBODY_RK4, 0.1, 3, sync=True)
            res.append(ctx.download_aos())
    for got in res:
        check_state(got.astype(np.float64), want, bodies, TOL64_TRAJ, steps=3)
    assert rel_inf(res[1][:, 1:4], res[0][:, 1:4]) < 1e-12


def test_sym_mode_can_be_switched_on_a_live_context(oracle):
    b = generators.uniform_bodies(2000, seed=8)
    want = oracle.rhs(b, threads=8)[:, 4:7]
    with capi.Context(2000, 64) as ctx:
        ctx.upload_aos(b)
        a0 = ctx.rhs()
        ctx.set_option("pair_symmetry", 1)
        a1 = ctx.rhs()
        ctx.set_option("pair_symmetry", 0)
        a2 = ctx.rhs()
    assert np.array_equal(a0, a2)                      # the default path is deterministic
    assert rel_inf(a1, want) < TOL64_RHS and rel_inf(a0, want) < TOL64_RHS


def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.skipif(_n_gpus() < 2, reason="needs >= 2 GPUs")
def test_sym_two_ranks_match_oracle():
    """Two ranks under torchrun (tests/tools/multi_check.py): every rank walks its blocks of the
    one global schedule and the j-side sums are reduce-scattered over the ranks (NCCL)."""
    import os
    import subprocess
    import sys
    from conftest import ROOT
    env = dict(os.environ, NBODY_P2P="0", NBODY_PAIR_SYM="1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                        "--nproc-per-node", "2", "--master-addr", "127.0.0.1", "--master-port",
                        "29535", os.path.join(ROOT, "tests", "tools", "multi_check.py"), "3000"],
                       capture_output=True, text=True, timeout=280, env=env)
    assert r.returncode == 0 and "MULTI-GPU CHECK PASSED" in r.stdout, \
        r.stdout[-3000:] + r.stderr[-3000:]

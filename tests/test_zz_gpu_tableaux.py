"""General explicit Runge-Kutta tableaux and embedded pairs on the GPU (nbody_step_rk,
nbody_step_embedded, nbody_step_undo) against the oracle's generic integrator
(oracle_rk_general_f64 over the reference RHS).  Not in the reference.  Written after the round's
GPU budget was spent -- the same cases run on the CPU through the emulated library
(tests/test_capi_emulation.py) -- so the file sorts last: a failure here must not mask the
default path's suite under `pytest -x`."""
import numpy as np
import pytest

from conftest import rel_inf
from parallel_comp_labs_b200 import capi, generators, tableaux
from test_gpu_parity import check_state

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300, method="thread")]


@pytest.mark.parametrize("name", sorted(tableaux.TABLEAUX))
def test_tableau_vs_oracle(oracle, name):
    t = tableaux.TABLEAUX[name]
    plu = generators.plummer_bodies(1500, seed=11)
    want, _ = oracle.rk_general(plu, t["a"], t["b"], 2e-3, 3, eps=1e-3, threads=8)
    with capi.Context(1500, 64) as ctx:
        ctx.set_params(6.67e-11, 1e-3)
        ctx.upload_aos(plu)
        ctx.step_rk(t["a"], t["b"], 2e-3, 3, sync=True)
        check_state(ctx.download_aos(), want, plu, 1e-10, steps=3)
        assert ctx.timing().force_launches == 3 * len(t["b"]) - (3 - 1) * tableaux.is_fsal(name)


def test_tableau_rk4_equals_hardwired_rk4_and_fp32(oracle):
    t = tableaux.TABLEAUX["rk4"]
    plu = generators.plummer_bodies(3000, seed=12)
    with capi.Context(3000, 64) as ctx:
        ctx.set_params(6.67e-11, 1e-3)
        ctx.upload_aos(plu)
        ctx.step(capi.NBODY_RK4, 2e-3, 3, sync=True)
        hard = ctx.download_aos()
        ctx.upload_aos(plu)
        ctx.step_rk(t["a"], t["b"], 2e-3, 3, sync=True)
        assert rel_inf(ctx.download_aos()[:, 1:7], hard[:, 1:7]) < 1e-13
    uni = generators.uniform_bodies(2000, seed=4)
    t = tableaux.TABLEAUX["kutta3"]
    want, _ = oracle.rk_general(uni, t["a"], t["b"], 0.1, 3, threads=8)
    with capi.Context(2000, 32) as ctx:
        ctx.upload_aos(uni)
        ctx.step_rk(t["a"], t["b"], 0.1, 3, sync=True)
        got = ctx.download_aos().astype(np.float64)
    assert rel_inf(got[:, 1:4], want[:, 1:4]) < 1e-4 and rel_inf(got[:, 4:7], want[:, 4:7]) < 1e-4


@pytest.mark.parametrize("name", ["bs23", "rkf45", "cashkarp", "dopri5"])
def test_embedded_step_norms_and_undo(oracle, name):
    t = tableaux.TABLEAUX[name]
    plu = generators.plummer_bodies(1200, seed=13)
    want, werr = oracle.rk_general(plu, t["a"], t["b"], 5e-3, 1, b_err=t["b_err"], eps=1e-3, threads=8)
    with capi.Context(1200, 64) as ctx:
        ctx.set_params(6.67e-11, 1e-3)
        ctx.upload_aos(plu)
        norms = ctx.step_embedded(t["a"], t["b"], t["b_err"], 5e-3)
        check_state(ctx.download_aos(), want, plu, 1e-10)
        assert np.allclose(norms[:2], werr[:2], rtol=1e-4, atol=0), (norms, werr)
        assert np.allclose(norms[2:], werr[2:], rtol=1e-12, atol=0)
        ctx.step_undo()
        assert np.array_equal(ctx.download_aos(), plu)
        with pytest.raises(capi.NBodyError):
            ctx.step_undo()


def test_force_evaluations_saved_by_fsal_and_after_undo(oracle):
    t = tableaux.TABLEAUX["dopri5"]
    plu = generators.plummer_bodies(1200, seed=14)
    want, _ = oracle.rk_general(plu, t["a"], t["b"], 2e-3, 1, eps=1e-3, threads=8)
    want, _ = oracle.rk_general(want, t["a"], t["b"], 1e-3, 1, eps=1e-3, threads=8)
    with capi.Context(1200, 64) as ctx:
        ctx.set_params(6.67e-11, 1e-3)
        ctx.upload_aos(plu)
        ctx.step_embedded(t["a"], t["b"], t["b_err"], 2e-3)
        assert ctx.timing().force_launches == 7
        ctx.step_embedded(t["a"], t["b"], t["b_err"], 4e-3)
        assert ctx.timing().force_launches == 6       # k_0 = last stage of the step before
        ctx.step_undo()
        ctx.step_embedded(t["a"], t["b"], t["b_err"], 1e-3)
        assert ctx.timing().force_launches == 6       # the retry reuses the rejected step's k_0
        check_state(ctx.download_aos(), want, plu, 1e-10, steps=2)
        ctx.set_params(6.67e-11, 2e-3)
        ctx.step_embedded(t["a"], t["b"], t["b_err"], 1e-3)
        assert ctx.timing().force_launches == 7


def test_adaptive_controller_on_the_reference_4body_problem(golden):
    """Step-size control over Dormand-Prince 5(4) against the accurate solution the reference
    ships (traj1.csv; fixture floor ~4e-6), and against what the reference's own fixed-step RK2
    reaches at its smallest step (2.8e-3)."""
    b4 = golden["body4"]
    with capi.Context(4, 64) as ctx:
        ctx.upload_aos(b4)
        info = tableaux.integrate_adaptive(ctx, "dopri5", 0.0, 20.0, 0.1, rtol=1e-10)
        fin = ctx.download_aos()
    assert abs(info["t"] - 20.0) < 1e-9
    assert np.abs(fin[0, 1:4] - golden["traj1"][200, 1:4]).max() < 2e-5

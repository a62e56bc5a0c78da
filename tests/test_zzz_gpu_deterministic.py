"""Option "deterministic" on the GPU: the pair-symmetric kernel with fixed-point row sums
(parallel_comp_labs_b200/csrc/fixed_sums.cuh).  The reference sums every row sequentially on one
thread (lab_4 struct_ode_system.hpp:60-75), so its results are reproducible bit for bit; the default
pair-symmetric kernel gives that up (floating-point RED, order = scheduling), this mode restores it:
same tolerances against the oracle, identical bits from run to run, context to context, replayed
graph or plain launches."""
import numpy as np
import pytest

from conftest import rel_inf
from parallel_comp_labs_b200 import capi, generators
from test_gpu_parity import check_state

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300, method="thread")]


def rhs(bodies, prec=64, eps=1e-3, **options):
    with capi.Context(bodies.shape[0], prec) as ctx:
        ctx.set_params(6.67e-11, eps)
        for k, v in options.items():
            ctx.set_option(k, v)
        ctx.upload_aos(bodies)
        outs = [ctx.rhs().astype(np.float64) for _ in range(3)]      # three launches on one context
        return outs, ctx.timing()


def per_row(got, want):
    return float((np.abs(got - want).max(axis=1) / np.maximum(np.abs(want).max(axis=1), 1e-300)).max())


@pytest.mark.parametrize("n,prec,general", [(3000, 64, False), (3000, 64, True), (20000, 64, True),
                                            (40000, 64, False), (3000, 32, False), (40000, 32, True)])
def test_deterministic_rhs_same_bits_every_run_and_within_tolerance(oracle, n, prec, general):
    b = generators.plummer_bodies(n, seed=41 + n)
    b[7, 1:4] = b[8, 1:4] + 1e-5                       # a pair inside the clamp radius
    b[n // 2, 1:4] = b[20, 1:4]                        # a coincident pair in different blocks
    if general:
        b[:, 0] *= np.linspace(0.5, 2.0, n)
        b[33, 0] = 0.0
    bb = b.astype(np.float32 if prec == 32 else np.float64)
    rows = slice(0, n) if n <= 3000 else slice(n // 2 - 256, n // 2 + 256)
    want = oracle.rhs(bb.astype(np.float64), rows.start, rows.stop, eps=1e-3, threads=16)[rows, 4:7]
    det, tm = rhs(bb, prec, pair_symmetry=1, deterministic=1)
    assert tm.pair_symmetry == 1 and tm.deterministic == 1
    assert tm.uniform_mass == (0 if general else 1)
    assert np.array_equal(det[0], det[1]) and np.array_equal(det[0], det[2])
    again, _ = rhs(bb, prec, pair_symmetry=1, deterministic=1)       # a second context
    assert np.array_equal(det[0], again[0])
    assert rel_inf(det[0][rows], want) < (1e-12 if prec == 64 else 1e-4)
    assert per_row(det[0][rows], want) < (1e-10 if prec == 64 else 2e-3)
    flt, tmf = rhs(bb, prec, pair_symmetry=1)
    assert tmf.deterministic == 0
    assert rel_inf(det[0], flt[0]) < (2e-13 if prec == 64 else 1e-4)


def test_deterministic_every_fixed_point_build(oracle):
    """Every fx entry of the variant tables (reachable as the twin of a floating-point entry)."""
    b = generators.plummer_bodies(5000, seed=3)
    want = oracle.rhs(b, eps=1e-3, threads=16)[:, 4:7]
    seen = set()
    for prec, nvar in ((64, 11), (32, 6)):
        bb = b.astype(np.float32 if prec == 32 else np.float64)
        for v in range(nvar):
            det, tm = rhs(bb, prec, pair_symmetry=1, sym_variant=v, deterministic=1)
            assert tm.deterministic == 1 and tm.sym_variant >= nvar
            seen.add((prec, tm.sym_variant))
            assert np.array_equal(det[0], det[1]) and np.array_equal(det[0], det[2])
            assert rel_inf(det[0], want) < (1e-12 if prec == 64 else 1e-4)
    assert len(seen) == 9                              # 5 FP64 + 4 FP32 fixed-point builds


@pytest.mark.parametrize("n,graph", [(777, 1), (3000, 0), (3000, 1), (20000, 1)])
def test_deterministic_rk4_states_are_bit_identical(oracle, n, graph):
    b = generators.uniform_bodies(n, seed=9)           # general masses, eps = 1e-6
    outs = []
    for g in (graph, 1 - graph, graph):
        with capi.Context(n, 64) as ctx:
            ctx.set_params(6.67e-11, 1e-6)
            ctx.set_option("pair_symmetry", 1)
            ctx.set_option("deterministic", 1)
            ctx.set_option("use_graph", g)
            ctx.upload_aos(b)
            ctx.step(capi.NBODY_RK4, 0.1, 3, sync=True)
            outs.append(ctx.download_aos())
            assert ctx.timing().deterministic == 1
    assert np.array_equal(outs[0], outs[1]) and np.array_equal(outs[0], outs[2])
    if n <= 3000:
        want, _ = oracle.rk4(b, 0.1, 0.0, 1.0, eps=1e-6, max_steps=3, rows=False, threads=16)
        check_state(outs[0].astype(np.float64), want, b, 1e-10)


def test_deterministic_refusals_and_poison():
    b = generators.uniform_bodies(4000, seed=2)
    with pytest.raises(capi.NBodyError):               # eps = 0: no bound on a contribution
        rhs(b, eps=0.0, pair_symmetry=1, deterministic=1)
    with pytest.raises(capi.NBodyError):               # dynamic range beyond the 192-bit window
        rhs(b, eps=1e-40, pair_symmetry=1, deterministic=1)
    with capi.Context(b.shape[0], 64) as ctx:          # the option must precede the upload
        ctx.upload_aos(b)
        with pytest.raises(capi.NBodyError):
            ctx.set_option("deterministic", 1)
    bad = b.copy()
    bad[100, 1] = np.nan
    flt, _ = rhs(bad, eps=1e-6, pair_symmetry=1)
    det, _ = rhs(bad, eps=1e-6, pair_symmetry=1, deterministic=1)
    assert np.array_equal(np.isnan(flt[0]).any(axis=1), np.isnan(det[0]).any(axis=1))
    _, tm = rhs(b, eps=1e-6, pair_symmetry=0)          # the one-sided kernel: fixed-order sums
    assert tm.deterministic == 1

"""nbody_rhs_noalloc through the C ABI, against nbody_rhs on the same bodies.

nbody_rhs_noalloc is the entry point for hosts that cannot let this library reserve device memory:
an application whose own pool already owns the whole GPU, a process with a fixed memory budget, a
loop that must not run into the device-wide synchronisation of cudaMalloc / cudaFree.  Such a host
passes device arrays it already has (masses, positions, output) and the kernel works in a fixed
array inside the library image -- which every call shares and which keeps the previous call's
contents.  So the property to test is not one call but SEVERAL IN A ROW with different sizes and
bodies: each must equal nbody_rhs (the context-based entry point, separately allocated memory) to
the tolerance the parity tests use for the RHS, 1e-12 relative in FP64 and 1e-4 in FP32 (both run
the pair-symmetric kernel, whose sums are floating-point atomics: equal up to rounding).

Here torch plays the host application that owns the device memory: the three arrays are torch
tensors, the library gets their addresses, and the work runs on torch's current (default) stream.

This file sorts after every other test file on purpose: `pytest -x` stops at the first failure,
and a failure here must not hide the rest of the GPU suite.  It passes alone in a fresh process
(the first call does the per-device set-up) as well as after the other tests."""
import numpy as np
import pytest

from conftest import rel_inf
from parallel_comp_labs_b200 import capi, generators

pytestmark = pytest.mark.gpu

TOL = {64: 1e-12, 32: 1e-4}


def context_rhs(bodies, prec, eps):
    """The existing entry point: nbody_create + upload + nbody_rhs, pair-symmetric kernel."""
    with capi.Context(bodies.shape[0], prec) as ctx:
        ctx.set_params(6.67e-11, eps)
        ctx.set_option("pair_symmetry", 1)
        ctx.upload_aos(bodies)
        return ctx.rhs().astype(np.float64)


def noalloc_rhs(torch, bodies, prec, eps):
    """The caller's side of nbody_rhs_noalloc: device arrays it owns, addresses in, sync, read."""
    n = bodies.shape[0]
    dt = torch.float64 if prec == 64 else torch.float32
    m = torch.tensor(bodies[:, 0], dtype=dt, device="cuda:0")
    r4 = torch.zeros((n, 4), dtype=dt, device="cuda:0")
    r4[:, :3] = torch.tensor(bodies[:, 1:4], dtype=dt, device="cuda:0")
    r4[:, 3] = 12345.0          # .w is documented as ignored
    acc = torch.full((n, 3), float("nan"), dtype=dt, device="cuda:0")
    capi.rhs_noalloc(n, m.data_ptr(), r4.data_ptr(), acc.data_ptr(), precision=prec, device=0,
                     eps=eps, stream=None)
    torch.cuda.synchronize()
    return acc.cpu().numpy().astype(np.float64)


@pytest.mark.parametrize("prec", [64, 32])
def test_noalloc_equals_context_rhs_on_calls_in_a_row(prec):
    torch = pytest.importorskip("torch")
    # sizes chosen so that every call finds the shared array dirty in a different way: a large
    # system first, then a ragged smaller one (stale rows beyond its padding), a tiny one, the
    # maximum, and the first one again
    cases = [(generators.uniform_bodies(5000, seed=11), 1e-6),
             (generators.plummer_bodies(1237, seed=12), 1e-3),
             (generators.uniform_bodies(3, seed=13), 1e-6),
             (generators.uniform_bodies(capi.NOALLOC_MAX_BODIES, seed=14), 1e-6),
             (generators.uniform_bodies(5000, seed=11), 1e-6)]
    first = None
    for k, (bodies, eps) in enumerate(cases):
        got = noalloc_rhs(torch, bodies, prec, eps)
        want = context_rhs(bodies, prec, eps)
        assert got.shape == want.shape and np.all(np.isfinite(got))
        assert rel_inf(got, want) < TOL[prec], (k, bodies.shape[0])
        if k == 0:
            first = got
    # the repeat of the first system, after three other calls used the array, is still right
    assert rel_inf(got, first) < TOL[prec]


def test_noalloc_two_back_to_back_calls_without_sync_between():
    """Two calls on ONE stream are ordered by the stream: no synchronisation between them is
    needed, the second simply starts from the array the first one left."""
    torch = pytest.importorskip("torch")
    a = generators.uniform_bodies(4096, seed=21)
    b = generators.uniform_bodies(2049, seed=22)
    outs = []
    keep = []
    for bodies in (a, b):
        n = bodies.shape[0]
        m = torch.tensor(bodies[:, 0], device="cuda:0")
        r4 = torch.zeros((n, 4), dtype=torch.float64, device="cuda:0")
        r4[:, :3] = torch.tensor(bodies[:, 1:4], device="cuda:0")
        acc = torch.empty((n, 3), dtype=torch.float64, device="cuda:0")
        capi.rhs_noalloc(n, m.data_ptr(), r4.data_ptr(), acc.data_ptr())
        keep.append((m, r4))
        outs.append(acc)
    torch.cuda.synchronize()
    for bodies, acc in zip((a, b), outs):
        assert rel_inf(acc.cpu().numpy(), context_rhs(bodies, 64, 1e-6)) < TOL[64]


def test_noalloc_rejects_sizes_beyond_the_array():
    with pytest.raises(capi.NBodyError, match="NBODY_NOALLOC_MAX_BODIES"):
        capi.rhs_noalloc(capi.NOALLOC_MAX_BODIES + 1, 1, 1, 1)
    capi.rhs_noalloc(0, 0, 0, 0)      # nothing to do is not an error

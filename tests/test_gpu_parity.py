"""Parity of the CUDA path (through the C ABI, libnbody_b200.so) against the oracle and the
committed golden fixtures.  Tolerances (BASELINE.json north_star): positions and velocities
within 1e-10 relative for FP64 and 1e-4 for FP32 over a short horizon; the RHS itself is held
to 1e-12 (FP64).  "Relative" = inf-norm of the difference over the inf-norm of the reference
quantity; velocities are compared through their INCREMENTS v(t)-v(0) as well, because on the
bundled random bodies free streaming hides force errors (SURVEY.md 4 warning)."""
import numpy as np
import pytest

from conftest import rel_inf
from parallel_comp_labs_b200 import capi, generators, lab

pytestmark = pytest.mark.gpu

TOL64_RHS, TOL64_TRAJ, TOL32 = 1e-12, 1e-10, 1e-4


def gpu_rhs(bodies, prec=64, eps=1e-6, G=6.67e-11, variant=-1, slices=0, sym=-1):
    """sym = -1: the library's automatic choice of force kernel (pair-symmetric from 2048 bodies
    on), 0: the one-sided kernel (what `variant` selects builds of), 1: the pair-symmetric one."""
    with capi.Context(bodies.shape[0], prec) as ctx:
        ctx.set_params(G, eps)
        ctx.set_option("variant", variant)
        ctx.set_option("slices", slices)
        ctx.set_option("pair_symmetry", 0 if (variant >= 0 or slices) and sym < 0 else sym)
        ctx.upload_aos(bodies)
        return ctx.rhs().astype(np.float64)


def gpu_run(bodies, scheme, tau, steps, prec=64, eps=1e-6, G=6.67e-11, graph=0, sym=-1):
    with capi.Context(bodies.shape[0], prec) as ctx:
        ctx.set_params(G, eps)
        ctx.set_option("use_graph", graph)
        ctx.set_option("pair_symmetry", sym)
        ctx.upload_aos(bodies)
        ctx.step(scheme, tau, steps, sync=True)
        return ctx.download_aos().astype(np.float64)


def check_state(got, want, init, tol, steps=1):
    assert np.array_equal(got[:, 0], want[:, 0])                      # masses untouched
    assert rel_inf(got[:, 1:4], want[:, 1:4]) < tol
    assert rel_inf(got[:, 4:7], want[:, 4:7]) < tol
    # velocity INCREMENTS (the force-sensitive quantity).  v(t)-v(0) cancels: its rounding floor
    # is ~ (4 roundings per step) * eps_mach * |v| / |dv|, e.g. the 4-body orbit returns to its
    # start after one period (|dv| ~ 1e-5 |v|); a wrong force gives an O(1) error here.
    dv_w = want[:, 4:7] - init[:, 4:7]
    if np.abs(dv_w).max() > 0:
        eps_mach = 2.2e-16 if tol < 1e-6 else 1.2e-7
        floor = 4 * steps * eps_mach * np.abs(want[:, 4:7]).max() / np.abs(dv_w).max()
        assert rel_inf(got[:, 4:7] - init[:, 4:7], dv_w) < max(tol, floor)


# ------------------------------------------------------------------ RHS
def test_rhs_golden_fixtures(golden, rand10k):
    assert rel_inf(gpu_rhs(golden["body4"]), golden["ref_rhs_body4"]) < TOL64_RHS
    assert rel_inf(gpu_rhs(golden["rand512"]), golden["ref_rhs_rand512"]) < TOL64_RHS
    uni = generators.uniform_bodies(777, seed=0x5EED0001)
    assert rel_inf(gpu_rhs(uni), golden["ref_rhs_uniform777"]) < TOL64_RHS
    plu = golden["plummer1000_in"]       # has a pair inside the clamp radius and a coincident pair
    got = gpu_rhs(plu, eps=1e-3)
    assert np.all(np.isfinite(got))
    assert rel_inf(got, golden["ref_rhs_plummer1000_eps1e-3"]) < TOL64_RHS
    # the clamped pair itself, relative to its own magnitude
    for i in (10, 11, 20, 21):
        assert rel_inf(got[i], golden["ref_rhs_plummer1000_eps1e-3"][i]) < 1e-11
    bodies, rows32, last8 = rand10k      # the reference's own 10 000-body input
    got = gpu_rhs(bodies)
    scale = np.abs(rows32).max()
    assert np.abs(got[:32] - rows32).max() / scale < TOL64_RHS
    assert np.abs(got[9992:] - last8).max() / scale < TOL64_RHS


@pytest.mark.parametrize("n", [1, 2, 3, 255, 256, 257, 511, 513, 1000, 4097])
def test_rhs_ragged_sizes_vs_oracle(oracle, n):
    bodies = generators.uniform_bodies(n, seed=100 + n)
    want = oracle.rhs(bodies, threads=4)[:, 4:7]
    got = gpu_rhs(bodies)
    assert got.shape == (n, 3) and np.all(np.isfinite(got))
    if n == 1:
        assert np.all(got == 0) and np.all(want == 0)     # only the self term: exactly 0
    else:
        assert rel_inf(got, want) < TOL64_RHS


def test_rhs_all_variants_and_slicings(oracle):
    bodies = generators.plummer_bodies(3000, seed=5)
    bodies[7, 1:4] = bodies[8, 1:4] + 1e-5
    want = oracle.rhs(bodies, eps=1e-3, threads=8)[:, 4:7]
    var = 0
    while True:
        try:
            got = gpu_rhs(bodies, eps=1e-3, variant=var)
        except capi.NBodyError:
            break
        assert rel_inf(got, want) < TOL64_RHS, var
        var += 1
    assert var >= 4
    for s in (1, 2, 5, 12):
        assert rel_inf(gpu_rhs(bodies, eps=1e-3, slices=s), want) < TOL64_RHS
    v32 = 0
    while True:
        try:
            got = gpu_rhs(bodies, prec=32, eps=1e-3, variant=v32)
        except capi.NBodyError:
            break
        # the planted pair sits 1e-5 apart: in FP32 its separation itself is rounded, so the
        # two rows it dominates are excluded from the FP32 tolerance
        mask = np.ones(3000, bool)
        mask[[7, 8]] = False
        assert np.abs(got[mask] - want[mask]).max() / np.abs(want[mask]).max() < TOL32, v32
        v32 += 1
    assert v32 >= 3


def test_rhs_edge_cases(oracle):
    b = generators.uniform_bodies(600, seed=3)
    b[5, 0] = 0.0                       # zero-mass body: feels forces, exerts none
    b[40, 1:4] = b[41, 1:4]             # coincident pair: clamp -> exactly zero pair term
    b[100, 1:4] = b[101, 1:4] + 3e-7    # inside eps=1e-6
    want = oracle.rhs(b)[:, 4:7]
    got = gpu_rhs(b)
    assert np.all(np.isfinite(got)) and rel_inf(got, want) < TOL64_RHS
    for i in (100, 101):
        assert rel_inf(got[i], want[i]) < 1e-11
    # linearity in G and in the masses (size-independent property)
    g2 = gpu_rhs(b, G=2 * 6.67e-11)
    assert rel_inf(g2, 2 * got) < 1e-15
    b2 = b.copy()
    b2[:, 0] *= 4.0
    assert rel_inf(gpu_rhs(b2), 4 * got) < 1e-15
    # large eps: everything inside the clamp radius -> a_i = -G sum m_j (r_i-r_j)/eps^3
    got = gpu_rhs(b, eps=1e4)
    want = oracle.rhs(b, eps=1e4)[:, 4:7]
    assert rel_inf(got, want) < TOL64_RHS


def test_rhs_eps_zero_and_extreme_scales(oracle):
    b = generators.uniform_bodies(300, seed=13)
    # eps = 0: the self pair is 0 * (m / 0) = NaN in the reference (struct_ode_system.hpp:69-72)
    assert np.all(np.isnan(oracle.rhs(b, eps=0.0)[:, 4:7]))
    assert np.all(np.isnan(gpu_rhs(b, eps=0.0)))
    # wide dynamic range of positions and masses (still finite r^2)
    c = b.copy()
    c[:, 1:4] *= 1e40
    c[:, 0] *= 1e60
    want = oracle.rhs(c, eps=1e30)[:, 4:7]
    assert np.all(np.isfinite(want)) and rel_inf(gpu_rhs(c, eps=1e30), want) < TOL64_RHS
    d = b.copy()
    d[:, 1:4] *= 1e-40
    want = oracle.rhs(d, eps=1e-50)[:, 4:7]
    assert np.all(np.isfinite(want)) and rel_inf(gpu_rhs(d, eps=1e-50), want) < TOL64_RHS


def test_rhs_large_n_row_subsample(oracle):
    """N = 262 144 Plummer: full parity is infeasible on CPU; the oracle evaluates 48 rows
    against all N (exactly the begin,end API).  Plus Newton's third law: sum m_i a_i = 0."""
    n = 262144
    bodies = generators.plummer_bodies(n, seed=0x5EED0003)
    got = gpu_rhs(bodies, eps=1e-3)
    scale = np.abs(got).max()
    for begin in (0, 131072 - 8, n - 16):
        want = oracle.rhs(bodies, begin, begin + 16, eps=1e-3, threads=8)[begin:begin + 16, 4:7]
        assert np.abs(got[begin:begin + 16] - want).max() / np.abs(want).max() < TOL64_RHS
        # per row, relative to the row's own magnitude (a halo body cannot hide behind the core)
        per_row = np.abs(got[begin:begin + 16] - want).max(axis=1) / np.abs(want).max(axis=1)
        assert per_row.max() < 1e-10
    mom = (bodies[:, 0:1] * got).sum(axis=0)
    assert np.abs(mom).max() / (bodies[:, 0].sum() * scale) < 1e-12


# ------------------------------------------------------------------ integrators
@pytest.mark.parametrize("tau,key,stride", [(0.1, "0100", 1), (0.05, "0050", 8), (0.025, "0025", 8)])
def test_rk2_4body_every_row_vs_reference(golden, tau, key, stride):
    """The reference's own RK2 trajectory of its 4-body problem, every stored row."""
    b4 = golden["body4"]
    ref = golden[f"ref_rk2_4body_{key}"]
    nsteps = lab.step_count(0.0, 20.0, tau)
    assert nsteps + 1 == int(golden[f"ref_rk2_4body_{key}_nrows"][0])
    rows = [np.concatenate([[0.0], b4[:, 1:4].ravel()])]
    with capi.Context(4, 64) as ctx:
        ctx.upload_aos(b4)
        t = 0.0 + tau
        for k in range(nsteps):
            ctx.step(capi.NBODY_RK2, tau, 1)
            rows.append(np.concatenate([[t], ctx.snapshot().ravel()]))
            t += tau
    rows = np.array(rows)[::stride]
    assert rows.shape == ref.shape
    assert np.array_equal(rows[:, 0], ref[:, 0])                        # time column, bit for bit
    assert np.abs(rows[:, 1:] - ref[:, 1:]).max() < TOL64_TRAJ          # |r| ~ 1
    assert np.all(rows[:, 3::3] == 0)                                   # z stays exactly 0


def test_rk4_vs_oracle_and_golden(oracle, golden):
    b4 = golden["body4"]
    got = gpu_run(b4, capi.NBODY_RK4, 0.1, 200, graph=1)
    check_state(got, golden["ref_rk4_4body_0100_final"], b4, TOL64_TRAJ, steps=200)
    uni = generators.uniform_bodies(777, seed=0x5EED0001)
    check_state(gpu_run(uni, capi.NBODY_RK4, 0.1, 3), golden["ref_rk4_uniform777_3steps"], uni,
                TOL64_TRAJ, steps=3)
    plu = golden["plummer1000_in"]
    check_state(gpu_run(plu, capi.NBODY_RK4, 1e-3, 3, eps=1e-3),
                golden["ref_rk4_plummer1000_3steps"], plu, TOL64_TRAJ, steps=3)


def test_rk2_rand512_vs_reference_rows(golden):
    b = golden["rand512"]
    with capi.Context(512, 64) as ctx:
        ctx.upload_aos(b)
        for k, idx in ((1, 0), (6, 1), (11, 2)):
            ctx.step(capi.NBODY_RK2, 0.1, k - (0 if idx == 0 else (1, 6)[idx - 1]), sync=True)
            pos = ctx.snapshot().ravel()
            ref = golden["ref_rk2_rand512_rows"][idx, 1:]
            assert np.abs(pos - ref).max() / np.abs(ref).max() < TOL64_TRAJ


def test_rk2_rk4_uniform20k_config(oracle):
    """BASELINE.json configs[1] (uniform N=20k FP64): 2 steps against the oracle."""
    b = generators.uniform_bodies(20000, seed=0x5EED0001)
    want, _ = oracle.rk4(b, 0.1, 0.0, 1.0, max_steps=2, rows=False, threads=oracle.max_threads())
    check_state(gpu_run(b, capi.NBODY_RK4, 0.1, 2), want, b, TOL64_TRAJ, steps=2)


def test_known_answer_order_of_accuracy(golden):
    """Against the reference's accurate solution traj{1..4}.csv."""
    truth = np.stack([golden[f"traj{k}"][:, 1:4] for k in (1, 2, 3, 4)], axis=1)
    errs = {}
    for scheme, tau, every in ((2, 0.1, 1), (2, 0.05, 2), (2, 0.025, 4), (4, 0.1, 1)):
        with capi.Context(4, 64) as ctx:
            ctx.upload_aos(golden["body4"])
            pos = []
            for k in range(200):
                ctx.step(scheme, tau, every)
                pos.append(ctx.snapshot())
        errs[(scheme, tau)] = np.abs(np.array(pos) - truth[1:]).max()
    assert abs(errs[(2, 0.1)] - 4.0853e-2) < 5e-6
    assert abs(errs[(2, 0.05)] - 1.0966e-2) < 5e-6
    assert abs(errs[(2, 0.025)] - 2.8310e-3) < 5e-6
    assert abs(errs[(4, 0.1)] - 1.2157e-5) < 5e-8


TABLEAUX = {  # name: (a_sub, b, scale)  -- explicit RK, stage s+1 depends on stage s only
    "heun2": ([1.0], [1.0, 1.0], 0.5),
    "ralston2": ([2.0 / 3.0], [1.0, 3.0], 0.25),
    "heun3": ([1.0 / 3.0, 2.0 / 3.0], [1.0, 0.0, 3.0], 0.25),
    "rk4_as_data": ([0.5, 0.5, 1.0], [1.0, 2.0, 2.0, 1.0], 1.0 / 6.0),
}


@pytest.mark.parametrize("name", list(TABLEAUX))
def test_step_tableau_vs_oracle(oracle, golden, name):
    """nbody_step_tableau (tableau as data; not in the reference) against the oracle's generic
    sub-diagonal integrator, and its order of accuracy on the reference's 4-body problem."""
    a, b, scale = TABLEAUX[name]
    bodies = generators.uniform_bodies(1200, seed=33)
    want = oracle.rk_subdiag(bodies, a, b, scale, 0.1, 0.0, 1.0, max_steps=3, threads=4)
    with capi.Context(1200, 64) as ctx:
        ctx.upload_aos(bodies)
        ctx.step_tableau(a, b, scale, 0.1, 3, sync=True)
        got = ctx.download_aos()
    check_state(got, want, bodies, TOL64_TRAJ, steps=3)
    truth = np.stack([golden[f"traj{k}"][:, 1:4] for k in (1, 2, 3, 4)], axis=1)
    with capi.Context(4, 64) as ctx:
        ctx.upload_aos(golden["body4"])
        pos = []
        for _ in range(200):
            ctx.step_tableau(a, b, scale, 0.1, 1)
            pos.append(ctx.snapshot())
    err = np.abs(np.array(pos) - truth[1:]).max()
    lo, hi = {"heun2": (5e-3, 2e-1), "ralston2": (5e-3, 2e-1), "heun3": (1e-4, 1e-2),
              "rk4_as_data": (5e-6, 5e-5)}[name]
    assert lo < err < hi, err


def test_graph_replay_equals_plain_launches(golden):
    b = generators.uniform_bodies(1500, seed=8)
    a = gpu_run(b, capi.NBODY_RK4, 0.05, 6, graph=0)
    c = gpu_run(b, capi.NBODY_RK4, 0.05, 6, graph=1)
    assert np.array_equal(a, c)


def test_uniform_mass_specialisation(oracle):
    """Equal masses are detected at upload and factored out of the pair sum (one FP64
    instruction less per pair); the result must agree with the general kernel and the oracle,
    including pad bodies (parked at 1e150 instead of m = 0) and clamped pairs."""
    for n in (4, 257, 3000):
        b = generators.plummer_bodies(n, seed=40 + n)
        if n > 100:
            b[7, 1:4] = b[8, 1:4] + 1e-5
            b[9, 1:4] = b[10, 1:4]
        want = oracle.rhs(b, eps=1e-3, threads=4)[:, 4:7]
        outs = {}
        for um in (-1, 0):
            with capi.Context(n, 64) as ctx:
                ctx.set_params(6.67e-11, 1e-3)
                ctx.set_option("unimass", um)
                ctx.upload_aos(b)
                outs[um] = ctx.rhs()
                ctx.step(capi.NBODY_RK4, 1e-3, 2, sync=True)
                outs[(um, "s")] = ctx.download_aos()
        assert np.all(np.isfinite(outs[-1]))
        assert rel_inf(outs[-1], want) < TOL64_RHS and rel_inf(outs[0], want) < TOL64_RHS
        assert rel_inf(outs[(-1, "s")][:, 1:7], outs[(0, "s")][:, 1:7]) < 1e-13
    b = generators.plummer_bodies(2000, seed=77).astype(np.float32)
    w32 = oracle.rhs(b.astype(np.float64), eps=1e-3, threads=4)[:, 4:7]
    assert rel_inf(gpu_rhs(b, prec=32, eps=1e-3), w32) < TOL32


def test_determinism():
    """pair_symmetry = 0 (the one-sided kernel) is bit-reproducible run to run: dynamic unit
    scheduling, but fixed-order slice reduction.  The automatic choice at this size is the
    pair-symmetric kernel, whose sums go through floating-point atomics: reproducible to rounding."""
    b = generators.plummer_bodies(5000, seed=2)
    a = gpu_run(b, capi.NBODY_RK4, 1e-3, 2, eps=1e-3, sym=0)
    c = gpu_run(b, capi.NBODY_RK4, 1e-3, 2, eps=1e-3, sym=0)
    assert np.array_equal(a, c)
    d = gpu_run(b, capi.NBODY_RK4, 1e-3, 2, eps=1e-3)
    e = gpu_run(b, capi.NBODY_RK4, 1e-3, 2, eps=1e-3)
    assert rel_inf(d[:, 1:], e[:, 1:]) < 1e-13 and rel_inf(d[:, 1:], a[:, 1:]) < 1e-12


# ------------------------------------------------------------------ FP32
def test_fp32_path(oracle, golden):
    b4 = golden["body4"]
    # lab_5's committed FP32 GPU trajectory, through lab_5's own entry point and SoA layout
    ref = golden["lab5_res4body_async"]
    r4 = np.zeros((4, 4), np.float32)
    v4 = np.zeros((4, 4), np.float32)
    r4[:, :3], v4[:, :3] = b4[:, 1:4], b4[:, 4:7]
    with capi.Context(4, 32) as ctx:
        ctx.upload_soa(b4[:, 0].astype(np.float32), r4, v4)
        rows = []
        for k in range(200):
            ctx.step(capi.NBODY_RK2, 0.1, 1)
            rows.append(ctx.snapshot().ravel())
    rows = np.array(rows, dtype=np.float64)
    assert np.abs(rows[0] - ref[0, 1:]).max() < 2e-7
    assert np.abs(rows - ref[:, 1:]).max() < TOL32 * 2
    # FP32 vs the FP64 oracle on a random system, 5 steps
    b = generators.uniform_bodies(2000, seed=4)
    want, _ = oracle.rk2(b, 0.1, 0.0, 1.0, max_steps=5, rows=False, threads=8)
    got = gpu_run(b, capi.NBODY_RK2, 0.1, 5, prec=32)
    assert rel_inf(got[:, 1:4], want[:, 1:4]) < TOL32
    assert rel_inf(got[:, 4:7], want[:, 4:7]) < TOL32
    a32 = gpu_rhs(b, prec=32)
    assert rel_inf(a32, oracle.rhs(b, threads=8)[:, 4:7]) < TOL32


# ------------------------------------------------------------------ boundary behaviour
def test_soa_upload_equals_aos(golden):
    b = generators.uniform_bodies(900, seed=6)
    r4 = np.zeros((900, 4))
    v4 = np.zeros((900, 4))
    r4[:, :3], v4[:, :3] = b[:, 1:4], b[:, 4:7]
    r4[:, 3] = 123.0                                   # T3.w is ignored
    with capi.Context(900, 64) as ctx:
        ctx.upload_soa(b[:, 0], r4, v4)
        ctx.step(capi.NBODY_RK2, 0.1, 2, sync=True)
        a = ctx.download_aos()
    assert np.array_equal(a, gpu_run(b, capi.NBODY_RK2, 0.1, 2))


def test_snapshot_ring_and_errors():
    b = generators.uniform_bodies(300, seed=9)
    with capi.Context(300, 64) as ctx:
        with pytest.raises(capi.NBodyError):
            ctx.step(capi.NBODY_RK2, 0.1, 1)          # before upload
        ctx.upload_aos(b)
        with pytest.raises(capi.NBodyError):
            ctx.step(3, 0.1, 1)                        # unknown scheme
        tickets = [ctx.snapshot_async() for _ in range(capi.SNAPSHOT_SLOTS)]
        with pytest.raises(capi.NBodyError):
            ctx.snapshot_async()                       # ring full
        first = ctx.snapshot_wait(tickets[0])
        assert np.array_equal(first, b[:, 1:4])
        for t in tickets:
            ctx.snapshot_release(t)
        ctx.step(capi.NBODY_RK2, 0.1, 1)
        t1 = ctx.snapshot_async()                      # taken while nothing is synchronised
        ctx.step(capi.NBODY_RK2, 0.1, 1)
        p1 = ctx.snapshot_wait(t1)
        ctx.sync()
        p2 = ctx.download_aos()[:, 1:4]
        assert not np.array_equal(p1, p2)
        assert rel_inf(p1 + 0.1 * b[:, 4:7], p2) < 1e-3  # one more step of (almost) free streaming
        tm = ctx.timing()
        assert tm.steps_done == 2 and tm.force_launches == 2 and tm.interactions == 2 * 300 * 300


def test_energy_vs_oracle(oracle):
    b = generators.plummer_bodies(4000, seed=12)
    with capi.Context(4000, 64) as ctx:
        ctx.set_params(6.67e-11, 1e-3)
        ctx.upload_aos(b)
        k0, u0 = ctx.energy()
        ctx.step(capi.NBODY_RK4, 1e-3, 20, sync=True)
        k1, u1 = ctx.energy()
        fin = ctx.download_aos()
    ko, uo = oracle.energy(b, eps=1e-3, threads=8)
    assert abs(k0 - ko) / abs(ko) < 1e-13 and abs(u0 - uo) / abs(uo) < 1e-12
    ko1, uo1 = oracle.energy(fin, eps=1e-3, threads=8)
    assert abs((k1 + u1) - (ko1 + uo1)) / abs(ko1 + uo1) < 1e-11
    assert abs(u0 / k0 + 2.0) < 0.15                          # virial equilibrium of the model
    assert abs((k1 + u1) - (k0 + u0)) / abs(k0 + u0) < 1e-6   # RK4 energy drift, 20 steps


def test_momentum_and_com_are_conserved():
    """Newton's third law at the integrator level: total momentum and the centre-of-mass
    drift are conserved by RK4 to round-off (size-independent property)."""
    b = generators.uniform_bodies(6000, seed=17)
    with capi.Context(6000, 64) as ctx:
        ctx.upload_aos(b)
        d0 = ctx.diagnostics()
        ctx.step(capi.NBODY_RK4, 0.1, 10, sync=True)
        d1 = ctx.diagnostics()
    m = b[:, 0]
    assert abs(d0["mass"] - m.sum()) / m.sum() < 1e-14
    assert np.abs(d0["momentum"] - (m[:, None] * b[:, 4:7]).sum(0)).max() / np.abs(m[:, None] * b[:, 4:7]).sum() < 1e-14
    assert np.abs(d0["com"] - (m[:, None] * b[:, 1:4]).sum(0) / m.sum()).max() < 1e-12
    scale = np.abs(m[:, None] * b[:, 4:7]).sum()
    assert np.abs(d1["momentum"] - d0["momentum"]).max() / scale < 1e-14
    v_com = d0["momentum"] / d0["mass"]
    assert np.abs(d1["com"] - (d0["com"] + v_com * 1.0)).max() < 1e-11


def test_python_host_mirror_writes_reference_files(golden, tmp_path):
    sys_ = lab.ode_system()
    sys_.name, sys_.tau, sys_.t_start, sys_.t_final = "4Body_test", 0.1, 0.0, 20.0
    sys_.n_time_steps = 200
    sys_.traj_init = golden["body4"]
    solver = lab.Solver_N_body(output_folder=str(tmp_path))
    solver.solve_parallel_CUDA(sys_)
    lines = open(tmp_path / "results_0.100000.csv").read().splitlines()
    assert lines[0] + "\n" == lab.LAB4_HEADER and len(lines) == 202
    rows = np.array([[float(x) for x in l.split(";")[:-1]] for l in lines[1:]])
    assert np.abs(rows - golden["ref_rk2_4body_0100"]).max() < TOL64_TRAJ
    tcsv = open(tmp_path / "output_4Body_test_0.100000_np_1.csv").read().splitlines()
    assert tcsv[0] == "size;t_start;t_finish;tau;time_steps;num_procs;time_taken;avg_time_for_step"
    assert tcsv[1].startswith("4;0;20;0.1;200;1;")


def test_graph_is_not_replayed_across_uploads(oracle):
    """Re-upload with a different equal mass on one context, stepping through the CUDA-graph path:
    the captured step bakes the uniform mass into K2's parameters, so the graph must be dropped;
    an upload that changes nothing the capture holds keeps it (the end-to-end loop's case)."""
    n = 64
    b = generators.uniform_bodies(n, seed=2)
    general = b[:, 0].copy()
    with capi.Context(n, 64) as ctx:
        ctx.set_option("use_graph", 1)
        # (masses, graph captures expected so far): the same state uploaded again keeps the graph,
        # another equal mass, distinct masses, and equal masses again each drop it
        for mass, captures in ((1e9, 1), (1e9, 1), (3e9, 2), (None, 3), (None, 3), (1e9, 4)):
            b[:, 0] = general if mass is None else mass
            want, _ = oracle.rk4(b, 0.1, 0.0, 1.0, max_steps=3, rows=False)
            ctx.upload_aos(b)
            ctx.step(capi.NBODY_RK4, 0.1, 3, sync=True)
            got = ctx.download_aos()
            tm = ctx.timing()
            assert tm.uniform_mass == (0 if mass is None else 1)
            assert tm.graph_captures == captures
            check_state(got.astype(np.float64), want, b, TOL64_TRAJ, steps=3)

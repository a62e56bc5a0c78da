"""The oracle (oracle/nbody_oracle.c) pinned against the reference: (1) bit-for-bit against
golden fixtures that tests/golden/make_golden.py generated from the reference's own code,
(2) bit-for-bit against oracle/_ref itself where that build exists, (3) against the
reference's own known-answer files (traj{1..4}.csv order-of-accuracy numbers of BASELINE.md 3,
res4body_async.csv)."""
import numpy as np
import pytest

from conftest import rel_inf


def test_rhs_bit_exact_vs_golden(oracle, golden, rand10k):
    assert np.array_equal(oracle.rhs(golden["body4"])[:, 4:7], golden["ref_rhs_body4"])
    assert np.array_equal(oracle.rhs(golden["rand512"])[:, 4:7], golden["ref_rhs_rand512"])
    from parallel_comp_labs_b200 import generators
    uni = generators.uniform_bodies(777, seed=0x5EED0001)
    assert np.array_equal(oracle.rhs(uni)[:, 4:7], golden["ref_rhs_uniform777"])
    plu = golden["plummer1000_in"]
    assert np.array_equal(oracle.rhs(plu, eps=1e-3)[:, 4:7], golden["ref_rhs_plummer1000_eps1e-3"])
    bodies, rows32, last8 = rand10k
    assert bodies.shape == (10000, 7)
    assert np.array_equal(oracle.rhs(bodies, 0, 32, threads=4)[:32, 4:7], rows32)
    assert np.array_equal(oracle.rhs(bodies, 9992, 10000)[9992:, 4:7], last8)


def test_rhs_semantics(oracle, golden):
    b = golden["body4"]
    d = oracle.rhs(b, 1, 3)
    assert np.array_equal(d[1:3, 1:4], b[1:3, 4:7])      # diff.r = traj.v
    assert np.all(d[:, 0] == 0)                           # diff.m never written
    assert np.all(d[0] == 0) and np.all(d[3] == 0)        # rows outside [begin,end) untouched
    # coincident bodies: the clamp makes the pair term exactly 0, no NaN
    c = b.copy()
    c[1, 1:4] = c[0, 1:4]
    assert np.all(np.isfinite(oracle.rhs(c)))


@pytest.mark.parametrize("key,tau,stride,nrows", [("0100", 0.1, 1, 201), ("0050", 0.05, 8, 401),
                                                   ("0025", 0.025, 8, 802)])
def test_rk2_rows_bit_exact_vs_golden(oracle, golden, key, tau, stride, nrows):
    _, rows = oracle.rk2(golden["body4"], tau, 0.0, 20.0)
    assert int(golden[f"ref_rk2_4body_{key}_nrows"][0]) == nrows == rows.shape[0] + 1
    ref = golden[f"ref_rk2_4body_{key}"]
    full = np.vstack([np.concatenate([[0.0], golden["body4"][:, 1:4].ravel()]), rows])
    assert np.array_equal(full[::stride], ref)


def test_rk2_rand512_vs_golden(oracle, golden):
    _, rows = oracle.rk2(golden["rand512"], 0.1, 0.0, 1.0)
    assert rows.shape[0] == 11 and int(golden["ref_rk2_rand512_nrows"][0]) == 12
    assert np.array_equal(rows[[0, 5, 10]], golden["ref_rk2_rand512_rows"])


def test_rk4_bit_exact_vs_golden(oracle, golden):
    from parallel_comp_labs_b200 import generators
    f, _ = oracle.rk4(golden["body4"], 0.1, 0.0, 20.0, rows=False)
    assert int(golden["ref_rk4_4body_0100_steps"][0]) == 200
    assert np.array_equal(f, golden["ref_rk4_4body_0100_final"])
    uni = generators.uniform_bodies(777, seed=0x5EED0001)
    f, _ = oracle.rk4(uni, 0.1, 0.0, 1.0, max_steps=3, rows=False)
    assert np.array_equal(f, golden["ref_rk4_uniform777_3steps"])
    f, _ = oracle.rk4(golden["plummer1000_in"], 1e-3, 0.0, 1.0, eps=1e-3, max_steps=3, rows=False)
    assert np.array_equal(f, golden["ref_rk4_plummer1000_3steps"])


def test_known_answer_trajectories(oracle, golden):
    """max abs error over 4 bodies x 200 samples vs the reference's own
    accurate solution traj{1..4}.csv: RK2 4.0853e-2 / 1.0966e-2 / 2.8310e-3 (order 2),
    RK4 1.2157e-5 at tau = 0.1."""
    truth = np.stack([golden[f"traj{k}"][:, 1:4] for k in (1, 2, 3, 4)], axis=1)  # (201,4,3)
    errs = {}
    for tau, every in ((0.1, 1), (0.05, 2), (0.025, 4)):
        _, rows = oracle.rk2(golden["body4"], tau, 0.0, 20.0)
        pos = rows[every - 1::every, 1:][:200].reshape(200, 4, 3)
        errs[tau] = np.abs(pos - truth[1:]).max()
    assert abs(errs[0.1] - 4.0853e-2) < 5e-6
    assert abs(errs[0.05] - 1.0966e-2) < 5e-6
    assert abs(errs[0.025] - 2.8310e-3) < 5e-6
    assert 3.5 < errs[0.1] / errs[0.05] < 4.1 and 3.5 < errs[0.05] / errs[0.025] < 4.1
    _, rows = oracle.rk4(golden["body4"], 0.1, 0.0, 20.0)
    e4 = np.abs(rows[:200, 1:].reshape(200, 4, 3) - truth[1:]).max()
    assert abs(e4 - 1.2157e-5) < 5e-8


def test_fp32_oracle_vs_reference_gpu_csv(oracle, golden):
    """lab_5's committed FP32 GPU trajectory (res4body_async.csv): the FP32 restatement
    follows it to FP32 round-off (~1e-6..1e-5)."""
    ref = golden["lab5_res4body_async"]
    assert ref.shape == (200, 13)
    _, rows = oracle.rk2_f32(golden["body4"].astype(np.float32), 0.1, 0.0, 20.0)
    assert rows.shape[0] == 200
    assert np.abs(rows[0].astype(np.float64) - ref[0]).max() < 2e-7
    assert np.abs(rows.astype(np.float64) - ref).max() < 2e-4
    assert np.array_equal(rows[:, 0].astype(np.float64), ref[:, 0])  # time column, accumulated in float


def test_step_counts(oracle):
    """SURVEY.md F4 probe table."""
    for (tau, T, n) in ((0.1, 1, 11), (0.1, 2, 20), (0.1, 20, 200), (0.05, 20, 400),
                        (0.025, 20, 801), (0.01, 1, 100)):
        assert oracle.step_count(0.0, T, tau) == n
    for (tau, T, n) in ((0.1, 1, 10), (0.1, 20, 200), (0.05, 20, 401), (0.025, 20, 801),
                        (0.01, 1, 101), (0.001, 1, 1001)):
        assert oracle.step_count_f32(0.0, T, tau) == n


def test_partition(oracle):
    assert oracle.partition(10, 3) == ([3, 3, 4], [0, 3, 6])
    assert oracle.partition(10000, 8) == ([1250] * 8, [1250 * i for i in range(8)])
    with pytest.raises(ValueError):
        oracle.partition(4, 5)


# ---- directly against the reference build (only where oracle/_ref exists) ---------------------
def test_oracle_vs_ref_build(oracle, reflab4, golden):
    from parallel_comp_labs_b200 import generators
    for bodies, eps in ((golden["body4"], 1e-6), (generators.uniform_bodies(300, 3), 1e-6),
                        (generators.plummer_bodies(400, 9), 1e-3)):
        assert np.array_equal(oracle.rhs(bodies, eps=eps), reflab4.rhs(bodies, eps=eps))
        f_o, _ = oracle.rk4(bodies, 0.01, 0.0, 1.0, eps=eps, max_steps=2, rows=False)
        f_r, _ = reflab4.rk4(bodies, 0.01, 0.0, 1.0, eps=eps, max_steps=2)
        assert np.array_equal(f_o, f_r)


def test_ref_threads_as_ranks(reflab4, golden, tmp_path):
    """The reference's own MPI decomposition on 1, 2 and 4 thread-ranks gives identical rows
    (each row is computed by exactly one rank; gathers only move data)."""
    from oracle import read_csv_rows
    out = []
    for p, bodyr in ((1, True), (2, True), (4, False)):
        path = str(tmp_path / f"r{p}.csv")
        reflab4.rk2_csv(golden["body4"], 0.1, 0.0, 2.0, path, n_ranks=p, use_bodyr=bodyr)
        out.append(read_csv_rows(path)[1])
    assert np.array_equal(out[0], out[1]) and np.array_equal(out[0], out[2])
    assert out[0].shape == (21, 13)
    with pytest.raises(RuntimeError):
        reflab4.rk2_csv(golden["body4"], 0.1, 0.0, 1.0, str(tmp_path / "x.csv"), n_ranks=5)

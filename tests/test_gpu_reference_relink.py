"""The reference's OWN host programs driving libnbody_b200.so.

oracle/_ref/relink_lab4 = the reference's unmodified lab_4 N_body.cpp + class_Solver_N_body.hpp +
struct_ode_system.hpp + read_write_utils.hpp compiled where they lie, with only the body of
ode_RK2_MPI replaced by the patch INTEGRATION.md shows (tests/integration/lab4_ode_RK2_MPI_b200.hpp).
oracle/_ref/relink_lab5 = the reference's unmodified lab_5 N_body.cu (its main, test_4body,
read_data, write_bodies_coords) with only ode_RK2_CUDA replaced
(tests/integration/lab5_ode_RK2_CUDA_b200.inc).  Both are built by oracle/Makefile in the container
that has /root/reference and travel to the GPU box as binaries; here they run on the B200 and the
files THEY write are compared with the reference-built golden rows."""
import os
import shutil
import subprocess

import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu

REF = os.path.join(ROOT, "oracle", "_ref")
BODY4 = os.path.join(ROOT, "tests", "golden", "_4body.txt")


def _rows(path):
    lines = open(path).read().splitlines()
    return lines[0], np.array([[float(x) for x in l.split(";") if x] for l in lines[1:]])


def _need(exe):
    p = os.path.join(REF, exe)
    if not os.path.exists(p):
        pytest.skip(f"{p} not built (needs /root/reference at build time)")
    return p


def test_lab4_program_relinked_default_test(golden, tmp_path):
    """N_body.cpp:69-79 (ode_system_default_test at tau = 0.1 / 0.05 / 0.025) through the
    reference's Solver_N_body::solve_parallel_MPI: file names, header, row counts, time column
    bit for bit, positions < 1e-10 against the reference build's own rows."""
    exe = _need("relink_lab4")
    src = tmp_path / "src"
    (src / "include" / "N_body" / "N_body_config").mkdir(parents=True)
    (tmp_path / "output").mkdir()
    shutil.copy(BODY4, src / "include" / "N_body" / "N_body_config" / "4body.txt")
    r = subprocess.run([exe, "default"], cwd=src, capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, NBODY_REF_SCRATCH=str(tmp_path)))
    assert r.returncode == 0, r.stdout + r.stderr
    for tau, key, stride in ((0.1, "0100", 1), (0.05, "0050", 8), (0.025, "0025", 8)):
        head, rows = _rows(tmp_path / "output" / f"results_{tau:.6f}.csv")
        assert head == "time;r11;r12;r13;r21;r22;r23;r31;r32;r33;r41;r42;r43"
        assert rows.shape[0] == int(golden[f"ref_rk2_4body_{key}_nrows"][0])
        ref = golden[f"ref_rk2_4body_{key}"]
        got = rows[::stride]
        assert got.shape == ref.shape
        assert np.array_equal(got[:, 0], ref[:, 0])
        assert np.abs(got[:, 1:] - ref[:, 1:]).max() < 1e-10
        timing = open(tmp_path / "output" / f"output_4Body_test_{tau:.6f}_np_1.csv").read().splitlines()
        assert timing[0] == "size;t_start;t_finish;tau;time_steps;num_procs;time_taken;avg_time_for_step"
        assert timing[1].startswith(f"4;0;20;{tau:g};")


def test_lab4_program_relinked_main_as_is(rand10k, tmp_path):
    """The reference's main() exactly as committed (N_body.cpp:38-87): ode_system_random_data on
    <output>/rand_bodies.txt (its own 10 000 bodies), solve_parallel_MPI(sys, "Server")."""
    exe = _need("relink_lab4")
    bodies, _, _ = rand10k
    (tmp_path / "src").mkdir()
    (tmp_path / "output").mkdir()
    with open(tmp_path / "output" / "rand_bodies.txt", "w") as f:
        f.write(f"{bodies.shape[0]}\n")
        for b in bodies:
            f.write("\t".join(repr(float(x)) for x in b) + "\n")
    r = subprocess.run([exe, "main"], cwd=tmp_path / "src", capture_output=True, text=True,
                       timeout=600, env=dict(os.environ, NBODY_REF_SCRATCH=str(tmp_path)))
    assert r.returncode == 0, r.stdout + r.stderr
    outs = [p for p in os.listdir(tmp_path / "output") if p.startswith("output_4Body_random")]
    assert len(outs) == 1, os.listdir(tmp_path / "output")
    lines = open(tmp_path / "output" / outs[0]).read().splitlines()
    assert lines[1].startswith("10000;0;1;0.1;10;1;")


def test_lab5_program_relinked_test_4body(tmp_path):
    """lab_5's main -> test_4body (N_body.cu:408-436,493-496): reads include/4body.txt, writes
    res4body_async.csv (header + one FP32 row per step) and the timing CSV; rows against lab_5's
    committed GPU output (tests/golden: res4body_async rows) to FP32 round-off."""
    exe = _need("relink_lab5")
    (tmp_path / "src" / "include").mkdir(parents=True)
    (tmp_path / "output").mkdir()
    shutil.copy(BODY4, tmp_path / "src" / "include" / "4body.txt")
    r = subprocess.run([exe], cwd=tmp_path / "src", capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, NBODY_REF5_SCRATCH=str(tmp_path)))
    assert r.returncode == 0, r.stdout + r.stderr
    head, rows = _rows(tmp_path / "output" / "res4body_async.csv")
    assert head == "time;" + ";".join(f"r_{i}_{j}" for i in range(1, 5) for j in (1, 2, 3))
    gold = dict(np.load(os.path.join(ROOT, "tests", "golden", "golden.npz")))
    ref = gold["lab5_res4body_async"]
    assert rows.shape == ref.shape
    assert np.array_equal(rows[:, 0], ref[:, 0])                 # FP32 time column, bit for bit
    assert np.abs(rows[:, 1:] - ref[:, 1:]).max() < 2e-4
    assert np.abs(rows[0, 1:] - ref[0, 1:]).max() < 1e-6         # first step: FP32 round-off
    timing = open(tmp_path / "output" / "res_4body_float_64_.csv").read().splitlines()
    assert timing[0] == "type;n_bodies;n_steps;block_size;time_taken;avg_exec_time"
    assert timing[1].startswith("float;4;200;64;")

"""The C++ host driver binary (parallel_comp_labs_b200/bin/nbody_b200) end to end: reads the
lab's body file, writes the lab's CSVs (both flavours), final-state dump is restartable."""
import json
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT
from parallel_comp_labs_b200 import generators, lab

pytestmark = pytest.mark.gpu
EXE = os.path.join(ROOT, "parallel_comp_labs_b200", "bin", "nbody_b200")
IN4 = os.path.join(ROOT, "tests", "golden", "_4body.txt")


def run(*args):
    r = subprocess.run([EXE, *map(str, args)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr + r.stdout
    return r.stdout


def parse(path):
    lines = open(path).read().splitlines()
    return lines[0], np.array([[float(x) for x in l.split(";")[:-1]] for l in lines[1:]])


def test_lab4_flavour_4body(golden, tmp_path):
    run("--system", "default_test", "--input", IN4, "--tau", 0.1, "--output-folder", tmp_path)
    header, rows = parse(tmp_path / "results_0.100000.csv")
    assert header + "\n" == lab.LAB4_HEADER
    ref = golden["ref_rk2_4body_0100"]
    assert rows.shape == ref.shape == (201, 13)
    assert np.array_equal(rows[:, 0], ref[:, 0])
    assert np.abs(rows - ref).max() < 1e-10
    raw = open(tmp_path / "results_0.100000.csv", "rb").read().splitlines()
    assert raw[1] == bytes(golden["ref_row_bytes"]).splitlines()[0].replace(
        b"3.0000000000000004e-01", b"0.0000000000000000e+00")        # t_start row, byte for byte
    t = open(tmp_path / "output_4Body_test_0.100000_np_1.csv").read().splitlines()
    assert t[0] == "size;t_start;t_finish;tau;time_steps;num_procs;time_taken;avg_time_for_step"
    assert t[1].startswith("4;0;20;0.1;200;1;")


def test_lab5_flavour_fp32(golden, tmp_path):
    run("--system", "default_test", "--input", IN4, "--precision", 32, "--flavour", "lab5",
        "--traj-out", tmp_path / "res4body_async.csv", "--timing-out", tmp_path / "t.csv")
    header, rows = parse(tmp_path / "res4body_async.csv")
    assert header + "\n" == lab.lab5_header(4)
    ref = golden["lab5_res4body_async"]
    assert rows.shape == ref.shape == (200, 13)                       # no t_start row
    assert np.array_equal(rows[:, 0], ref[:, 0])                      # time printed as float
    assert np.abs(rows - ref).max() < 2e-4
    t = open(tmp_path / "t.csv").read().splitlines()
    assert t[0] == "type;n_bodies;n_steps;block_size;time_taken;avg_exec_time"
    assert t[1].startswith("float;4;200;")


def test_random_system_restart_and_report(oracle, tmp_path):
    b = generators.uniform_bodies(1500, seed=21)
    inp = tmp_path / "rand_bodies.txt"
    with open(inp, "w") as f:
        f.write("1500\n")
        for r in b:
            f.write("\t".join("%.17g" % x for x in r) + "\n")
    out = run("--input", inp, "--scheme", "rk4", "--tau", 0.1, "--t-final", 0.35,
              "--final-state", tmp_path / "final.txt", "--timing-out", tmp_path / "t.csv",
              "--traj-out", tmp_path / "traj.csv", "--out-every", 2, "--report")
    rep = json.loads(out.strip().splitlines()[-1])
    steps = oracle.step_count(0.0, 0.35, 0.1)
    assert rep["steps"] == steps == 4 and rep["n"] == 1500
    fin = lab.read_data(str(tmp_path / "final.txt"))
    want, _ = oracle.rk4(b, 0.1, 0.0, 0.35, rows=False, threads=4)
    assert np.abs(fin[:, 1:4] - want[:, 1:4]).max() / np.abs(want[:, 1:4]).max() < 1e-10
    assert np.abs(fin[:, 4:7] - want[:, 4:7]).max() / np.abs(want[:, 4:7]).max() < 1e-10
    header, rows = parse(tmp_path / "traj.csv")
    assert rows.shape == (1 + 2, 1 + 3 * 1500)                        # t_start row + every 2nd step
    assert rep["energy_drift"] < 1e-3


def test_lab5_sweep_mode(tmp_path):
    """lab_5's test_random_bodies() (N_body.cu:438-491) as a driver mode: every <name>_bodies.txt
    present is run with RK2, T=[0,2], tau=0.1 and gets the reference's timing CSV."""
    inc, out = tmp_path / "include", tmp_path / "output"
    inc.mkdir()
    out.mkdir()
    assert subprocess.run([EXE, "--generate", "uniform", "--n", "10000", "--generate-out",
                           str(inc / "10k_bodies.txt")]).returncode == 0
    r = subprocess.run([EXE, "--lab5-sweep", str(inc), "--output-folder", str(out)],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "1 input file(s) run" in r.stdout, r.stdout + r.stderr
    t = open(out / "res_10k_float_128_.csv").read().splitlines()
    assert t[0] == "type;n_bodies;n_steps;block_size;time_taken;avg_exec_time"
    assert t[1].startswith("float;10000;20;")   # (2-0)/0.1 in float = 20


def test_per_body_files_and_binary_sidecar(golden, tmp_path):
    """The lab's 4-body run to t = 20 with RK4: every body's traj<k>.csv against the reference's own
    per-body fixture N_body_config/traj<k>.csv (max error 1.2157e-5 over 4 bodies x 200
    samples); the binary sidecar converts to the direct CSV and per-body files byte for byte; a
    20k-body run writes a range of bodies and a sidecar of 3 n sizeof(T) bytes per row."""
    d = tmp_path / "pb"
    d.mkdir()
    csv, binf = tmp_path / "traj.csv", tmp_path / "traj.bin"
    run("--system", "default_test", "--input", IN4, "--tau", 0.1, "--scheme", "rk4", "--traj-out", csv,
        "--traj-bin", binf, "--per-body-dir", d, "--timing-out", tmp_path / "t.csv")
    _, whole = parse(csv)
    assert whole.shape == (201, 13)
    worst = 0.0
    for k in range(4):
        rows = lab.read_per_body(str(d / f"traj{k + 1}.csv"))
        fix = golden[f"traj{k + 1}"]
        assert rows.shape == (201, 4) and np.array_equal(rows[:, 1:], whole[:, 1 + 3 * k:4 + 3 * k])
        worst = max(worst, float(np.abs(rows[:, 1:] - fix[:201, 1:4]).max()))
    assert abs(worst - 1.2157e-5) < 2e-8                           # the reference's known answer for RK4
    run("--bin2csv", binf, "--bin2csv-out", tmp_path / "again.csv")
    assert open(tmp_path / "again.csv", "rb").read() == open(csv, "rb").read()
    d2 = tmp_path / "pb2"
    d2.mkdir()
    run("--bin2perbody", binf, "--per-body-dir", d2)
    for k in range(1, 5):
        assert open(d2 / f"traj{k}.csv", "rb").read() == open(d / f"traj{k}.csv", "rb").read()
    t, pos, flavour = lab.read_traj_bin(str(binf))
    assert flavour == "lab4" and np.array_equal(pos.reshape(201, 12), whole[:, 1:])

    n = 20000
    inp = tmp_path / "in.txt"
    run("--generate", "uniform", "--n", n, "--generate-out", inp)
    d3 = tmp_path / "pb3"
    d3.mkdir()
    run("--input", inp, "--scheme", "rk4", "--tau", 0.1, "--t-final", 0.35, "--flavour", "lab5", "--traj-bin",
        tmp_path / "big.bin", "--per-body-dir", d3, "--per-body-first", 19999, "--timing-out", tmp_path / "t.csv")
    assert sorted(os.listdir(d3)) == ["traj19999.csv", "traj20000.csv"]
    t, pos, flavour = lab.read_traj_bin(str(tmp_path / "big.bin"))
    assert flavour == "lab5" and pos.shape == (4, n, 3)
    assert os.path.getsize(tmp_path / "big.bin") == 32 + 4 * (8 + 3 * n * 8)
    rows = lab.read_per_body(str(d3 / "traj20000.csv"))
    assert np.array_equal(rows[:, 1:], pos[:, n - 1, :]) and np.array_equal(rows[:, 0], t)


def test_convergence_study_reproduces_the_known_answers(golden, tmp_path):
    """`nbody_b200 --convergence DIR` = the reference's commented convergence runs (N_body.cpp:69-79)
    as one command: RK2 at tau = 0.1 / 0.05 / 0.025 against traj{1..4}.csv gives 4.0853e-2 /
    1.0966e-2 / 2.8310e-3 (order 2), RK4 at 0.1 gives 1.2157e-5."""
    for k in range(1, 5):
        with open(tmp_path / f"traj{k}.csv", "w") as f:
            f.write("time;r1;r2;r3\n")
            for row in golden[f"traj{k}"]:
                f.write(";".join(repr(float(x)) for x in row) + "\n")
    out = run("--convergence", tmp_path, "--input", IN4, "--output-folder", tmp_path)
    res = json.loads(out.strip().splitlines()[-1])
    assert res["taus"] == [0.1, 0.05, 0.025]
    for got, want in zip(res["max_abs_err"], (4.0853e-2, 1.0966e-2, 2.8310e-3)):
        assert abs(got - want) < 5e-6
    assert all(abs(o - 2.0) < 0.15 for o in res["observed_order"])
    out = run("--convergence", tmp_path, "--input", IN4, "--scheme", "rk4", "--taus", "0.1",
              "--output-folder", tmp_path)
    res = json.loads(out.strip().splitlines()[-1])
    assert abs(res["max_abs_err"][0] - 1.2157e-5) < 5e-8 and res["observed_order"] == []

"""C++ host driver (bin/nbody_b200) on 1..G GPUs: multi-GPU parity of the final state and the
cost of asynchronous trajectory output (FP32, N per GPU fixed, output
on vs off).  usage: driver_check.py N_PER_GPU GPUS [precision]"""
import json, os, subprocess, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from parallel_comp_labs_b200 import generators, lab
EXE = os.path.join(ROOT, "parallel_comp_labs_b200", "bin", "nbody_b200")
n_per, gpus = int(sys.argv[1]), int(sys.argv[2])
prec = int(sys.argv[3]) if len(sys.argv) > 3 else 32
tmp = "/dev/shm/nbody_driver_check"
os.makedirs(tmp, exist_ok=True)

def write_bodies(path, b):
    with open(path, "w") as f:
        f.write(f"{b.shape[0]}\n")
        np.savetxt(f, b, fmt="%.17g", delimiter="\t")

def run(inp, g, steps, traj, final=None):
    cmd = [EXE, "--input", inp, "--precision", str(prec), "--scheme", "rk4", "--tau", "1e-3",
           "--eps", "1e-3", "--gpus", str(g), "--max-steps", str(steps), "--t-final", "1e9",
           "--timing-out", f"{tmp}/t.csv", "--report", "--flavour", "lab5"]
    if traj: cmd += ["--traj-out", f"{tmp}/traj.csv"]
    if final: cmd += ["--final-state", final]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=1200)
    assert r.returncode == 0, r.stderr + r.stdout
    return json.loads(r.stdout.strip().splitlines()[-1])

res = {}
for g in sorted({1, gpus}):
    n = n_per * g
    inp = f"{tmp}/in_{n}.txt"
    t0 = time.time()
    subprocess.run([EXE, "--generate", "plummer", "--n", str(n), "--seed", "0x5EED0005",
                    "--generate-out", inp], check=True)   # C++ generator: seconds even at 4M bodies
    tw = time.time() - t0
    off = run(inp, g, 12, False)
    on = run(inp, g, 12, True)
    size = os.path.getsize(f"{tmp}/traj.csv")
    res[g] = {"n": n, "off_wall_s": off["wall_s"], "on_wall_s": on["wall_s"],
              "off_loop_s": off["loop_s"], "on_loop_s": on["loop_s"],
              "step_time_on_over_off": on["loop_s"] / off["loop_s"],
              "on_over_off": on["wall_s"] / off["wall_s"], "G_int_s_off": off["G_interactions_per_s"],
              "G_int_s_on": on["G_interactions_per_s"], "writer_busy_s": on["writer_busy_s"],
              "traj_bytes": size, "energy_drift": off["energy_drift"]}
    print(f"gpus={g} n={n}: loop off {off['loop_s']:.3f}s on {on['loop_s']:.3f}s ratio {on['loop_s']/off['loop_s']:.4f}; "
          f"wall off {off['wall_s']:.3f}s on {on['wall_s']:.3f}s ratio {on['wall_s']/off['wall_s']:.3f} "
          f"rate {off['G_interactions_per_s']:.0f} G/s writer busy {on['writer_busy_s']:.2f}s traj {size/1e6:.0f} MB", flush=True)
# multi-GPU parity of the driver: same input on 1 and on G GPUs
if gpus > 1:
    n = 20000
    inp = f"{tmp}/in_par.txt"
    write_bodies(inp, generators.uniform_bodies(n, seed=5))
    run(inp, 1, 3, False, f"{tmp}/f1.txt"); run(inp, gpus, 3, False, f"{tmp}/fg.txt")
    a, b = lab.read_data(f"{tmp}/f1.txt"), lab.read_data(f"{tmp}/fg.txt")
    err = np.abs(a - b).max() / np.abs(a).max()
    res["parity_1_vs_G"] = float(err)
    print("driver 1 GPU vs", gpus, "GPUs final-state rel diff", err)
    assert err < (1e-12 if prec == 64 else 1e-5)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", f"driver_check_{gpus}gpu_fp{prec}.json"), "w"), indent=1)

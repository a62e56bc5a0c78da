"""Python mirror of the lab's host interface over the C ABI (same names, argument meaning
and error behaviour as the reference), used by the parity tests and bench.py.  The
production host driver is the C++ one in ``parallel_comp_labs_b200/host`` -- this module
follows it function for function.

Reference interfaces mirrored (file:line relative to the reference root):
  read_data / write_bodies_coords   lab_4 .../utils/read_write_utils.hpp:12-52
  ode_system<T> + factories         lab_4 .../struct_ode_system.hpp:20-134
  Solver_N_body<T>::solve_parallel_MPI   lab_4 .../class_Solver_N_body.hpp:87-208
  ode_RK2_MPI loop control          lab_4 .../schemes/ode_RK2_MPI.hpp:65,72,98-99
  ode_RK2_CUDA                      lab_5 N_body.cu:208-406
All arithmetic of the path runs on the GPU through :mod:`.capi`; nothing here computes
forces or integrates.
"""
from __future__ import annotations

import dataclasses
import os

import numpy as np

from . import capi
from .capi import NBODY_RK2, NBODY_RK4


def read_data(path: str, dtype=np.float64) -> np.ndarray:
    """``N`` then N records ``m x y z vx vy vz``, any whitespace (tabs, CRLF) between
    tokens, parsed the way ``operator>>`` does.  Raises ``IOError`` where the reference
    prints "file is't open" and throws 1 (read_write_utils.hpp:15-19)."""
    try:
        with open(path, "rb") as f:
            tok = f.read().split()
    except OSError as e:
        print("file is't open")
        raise IOError(1) from e
    n = int(tok[0])
    want = 7 * n
    vals = np.array(tok[1:1 + want], dtype=np.float64)
    if vals.size == want:
        return vals.reshape(n, 7).astype(dtype)
    # short file: operator>> assigns field by field; at end of input the failing extraction
    # leaves its field alone and nothing is assigned any more -- the partly read record keeps
    # the previous body's remaining fields and every later record repeats it
    out = np.zeros((n, 7))
    k = vals.size
    out.reshape(-1)[:k] = vals
    rec, fld = divmod(k, 7)
    if rec < n:
        if rec > 0:
            out[rec, fld:] = out[rec - 1, fld:]
        out[rec + 1:] = out[rec]
    return out.astype(dtype)


def format_row(time: float, positions: np.ndarray) -> str:
    """One trajectory row: ``t;x1;y1;z1;...;zN;`` scientific, precision 16, trailing ';'
    (write_bodies_coords, read_write_utils.hpp:42-52; lab_5 N_body.cu:67-76)."""
    flat = np.asarray(positions, dtype=np.float64).reshape(-1)
    return "%.16e;" % float(time) + "".join("%.16e;" % x for x in flat) + "\n"


def write_bodies_coords(fout, time: float, positions: np.ndarray) -> None:
    fout.write(format_row(time, positions))
    fout.flush()  # std::endl


LAB4_HEADER = "time;r11;r12;r13;r21;r22;r23;r31;r32;r33;r41;r42;r43\n"  # ode_RK2_MPI.hpp:52


def lab5_header(n: int) -> str:
    """N_body.cu:236-246."""
    cols = [f"r_{i}_{j}" for i in range(1, n + 1) for j in (1, 2, 3)]
    return "time;" + ";".join(cols) + "\n"


# ---- the binary trajectory sidecar of `nbody_b200 --traj-bin` (host/nbody_host.hpp: TrajBinHeader)
TRAJ_BIN_MAGIC = b"NBTRJ01\0"


def read_traj_bin(path: str):
    """Returns (times[rows], positions[rows, n, 3], flavour) of a sidecar file; a truncated last
    record is dropped, as the C++ converter does."""
    with open(path, "rb") as f:
        head = f.read(32)
        if len(head) != 32 or head[:8] != TRAJ_BIN_MAGIC:
            raise ValueError("not a trajectory sidecar written by --traj-bin")
        n = int(np.frombuffer(head, dtype="<u8", count=1, offset=8)[0])
        elem, flavour = (int(x) for x in np.frombuffer(head, dtype="<u4", count=2, offset=16))
        if elem not in (4, 8) or flavour not in (4, 5) or n == 0:
            raise ValueError("not a trajectory sidecar written by --traj-bin")
        body = f.read()
    rec = 8 + 3 * n * elem
    rows = len(body) // rec
    dt = np.dtype([("t", "<f8"), ("r", "<f8" if elem == 8 else "<f4", (n, 3))])
    a = np.frombuffer(body, dtype=dt, count=rows)
    return a["t"].copy(), a["r"].copy(), "lab5" if flavour == 5 else "lab4"


def traj_bin_to_csv(bin_path: str, csv_path: str) -> int:
    """The whole-system CSV a direct run writes (same header, same rows)."""
    t, r, flavour = read_traj_bin(bin_path)
    with open(csv_path, "w") as f:
        f.write(lab5_header(r.shape[1]) if flavour == "lab5" else LAB4_HEADER)
        for k in range(t.size):
            f.write(format_row(t[k], r[k]))
    return int(t.size)


def format_per_body_row(time: float, xyz) -> str:
    """One row of a per-body file ``traj<k>.csv``: ``t;x;y;z`` (header ``time;r1;r2;r3``, the
    layout of the reference's fixtures N_body_config/traj1.csv:1)."""
    return "%.16e;%.16e;%.16e;%.16e\n" % (float(time), float(xyz[0]), float(xyz[1]), float(xyz[2]))


def read_per_body(path: str) -> np.ndarray:
    """rows x (t, x, y, z) of a per-body file -- ours or the reference's fixtures (which write
    ``0.`` and ``1.`` for whole numbers)."""
    with open(path) as f:
        lines = f.read().split()
    assert lines[0] == "time;r1;r2;r3"
    return np.array([[float(x) for x in ln.split(";")] for ln in lines[1:]])


def step_count(t_start, t_final, tau, dtype=np.float64) -> int:
    """Steps the reference's loop executes, t accumulated in T."""
    T = np.dtype(dtype).type
    t = T(t_start) + T(tau)
    steps = 0
    while True:
        steps += 1
        if t > T(t_final):
            break
        t = T(t + T(tau))
    return steps


@dataclasses.dataclass
class ode_system:
    """Problem descriptor (struct_ode_system.hpp:20-32)."""
    tau: float = 0.1
    t_start: float = 0.0
    t_final: float = 1.0
    n_time_steps: int = 10
    N_bodies: int = 0
    traj_init: np.ndarray | None = None
    name: str = ""
    eps: float = 1e-6
    G: float = 6.67e-11

    def ode_system_default_test(self, time_step=0.1, eps=1e-6, G=6.67e-11,
                                path="./include/N_body/N_body_config/4body.txt"):
        """struct_ode_system.hpp:34-81 (4-body known-answer problem, T=[0,20])."""
        self.name = "4Body_test"
        self.t_start, self.t_final, self.tau = 0.0, 20.0, time_step
        self.n_time_steps = int((self.t_final - self.t_start) / self.tau)
        self.traj_init = read_data(path)
        self.N_bodies = self.traj_init.shape[0]
        self.eps, self.G = eps, G
        return self

    def ode_system_random_data(self, N_of_bodies=100, time_step=0.1, eps=1e-6, G=6.67e-11,
                               path=None, bodies=None):
        """struct_ode_system.hpp:84-133 (T=[0,1]); ``N_of_bodies`` is ignored as in the
        reference -- the size comes from the file (or from ``bodies``)."""
        self.name = "4Body_random"
        self.t_start, self.t_final, self.tau = 0.0, 1.0, time_step
        self.n_time_steps = int((self.t_final - self.t_start) / self.tau)
        self.traj_init = np.asarray(bodies, dtype=np.float64) if bodies is not None \
            else read_data(path)
        self.N_bodies = self.traj_init.shape[0]
        self.eps, self.G = eps, G
        return self


class Solver_N_body:
    """Run harness (class_Solver_N_body.hpp:43-243) with the CUDA back end."""

    def __init__(self, tol=1e-6, output_folder="../output/", file_name=""):
        self.tol, self.output_folder, self.file_name = tol, output_folder, file_name
        self.notifications = False
        self.timer = 0.0
        self.iter = 0
        self.final_state = None
        self.last_timing = None

    def solve_parallel_CUDA(self, sys: ode_system, OS="PC", scheme=NBODY_RK2, precision=64,
                            rank=0, world=1, nccl_uid=None, device=0, traj_path=None,
                            max_steps=0):
        """Counterpart of solve_parallel_MPI(sys, OS).  Writes the timing CSV (rank 0),
        and for the "4Body_test" system the trajectory CSV ``results_<tau>.csv``
        (class_Solver_N_body.hpp:185-190), unless ``traj_path`` overrides it."""
        import time as _time
        n = sys.traj_init.shape[0]
        if world > n:  # class_Solver_N_body.hpp:115-117
            raise ValueError("The number of processes is greater than the problem dimension")
        dtype = np.float64 if precision == 64 else np.float32
        if self.notifications:
            print(f"{self.file_name}: solving... \t", end="")
        out_csv = None
        if rank == 0:
            os.makedirs(self.output_folder, exist_ok=True)
            out_csv = open(os.path.join(
                self.output_folder, "output_%s_%s_np_%d.csv" % (sys.name, "%f" % sys.tau, world)), "w")
            out_csv.write("size;t_start;t_finish;tau;time_steps;num_procs;time_taken;"
                          "avg_time_for_step\n")
        if traj_path is None and sys.name == "4Body_test":
            traj_path = os.path.join(self.output_folder, "results_%s.csv" % ("%f" % sys.tau))
        fout = open(traj_path, "w") if (traj_path and rank == 0) else None
        T = np.dtype(dtype).type
        with capi.Context(n, precision, device, rank, world, nccl_uid) as ctx:
            ctx.set_params(sys.G, sys.eps)
            ctx.upload_aos(sys.traj_init.astype(dtype))
            if fout:
                fout.write(LAB4_HEADER)
                write_bodies_coords(fout, sys.t_start, sys.traj_init[:, 1:4].astype(dtype))
            t = T(sys.t_start) + T(sys.tau)
            t0 = _time.perf_counter()
            steps = 0
            pending = None
            while True:  # ode_RK2_MPI.hpp:72-100
                ctx.step(scheme, float(T(sys.tau)), 1)
                steps += 1
                if fout:
                    ticket = ctx.snapshot_async()
                    if pending is not None:  # format row k-1 while step k runs
                        write_bodies_coords(fout, pending[0], ctx.snapshot_wait(pending[1]))
                        ctx.snapshot_release(pending[1])
                    pending = (float(t), ticket)
                if max_steps and steps >= max_steps:
                    break
                if t > T(sys.t_final):
                    break
                t = T(t + T(sys.tau))
            if pending is not None:
                write_bodies_coords(fout, pending[0], ctx.snapshot_wait(pending[1]))
                ctx.snapshot_release(pending[1])
            ctx.sync()
            self.timer = _time.perf_counter() - t0
            self.iter = steps
            self.final_state = ctx.download_aos()
            self.last_timing = ctx.timing()
        if fout:
            fout.close()
        if out_csv:
            out_csv.write("%d;%g;%g;%g;%d;%d;%g;%g\n" % (
                n, sys.t_start, sys.t_final, sys.tau, sys.n_time_steps, world, self.timer,
                self.timer / max(sys.n_time_steps, 1)))
            out_csv.close()
        if self.notifications:
            print("  Done!")
        return self.final_state


def ode_RK2_CUDA(m, r, v, t_start, t_final, tau, path="", path_1="", precision=32,
                 scheme=NBODY_RK2, eps=1e-6, G=6.67e-11, device=0):
    """lab_5's entry point (N_body.cu:208-406): SoA inputs (T m[n], T3 r[n], T3 v[n]),
    timing CSV at ``path``, trajectory CSV at ``path_1`` (N-body header, no t_start row,
    time printed as T).  Returns the final (n,7) state."""
    dtype = np.float64 if precision == 64 else np.float32
    T = np.dtype(dtype).type
    m = np.asarray(m, dtype=dtype)
    n = m.shape[0]
    fout = open(path, "w") if path else None
    if fout:
        fout.write("type;n_bodies;n_steps;block_size;time_taken;avg_exec_time\n")
    fout_1 = open(path_1, "w") if path_1 else None
    if fout_1:
        fout_1.write(lab5_header(n))
    with capi.Context(n, precision, device) as ctx:
        ctx.set_params(G, eps)
        ctx.upload_soa(m, r, v)
        t = T(t_start) + T(tau)
        total_ms = 0.0
        while True:
            ctx.step(scheme, float(T(tau)), 1)
            if fout_1:
                write_bodies_coords(fout_1, float(t), ctx.snapshot())
            total_ms += ctx.timing().last_call_ms
            if t > T(t_final):
                break
            t = T(t + T(tau))
        ctx.sync()
        final = ctx.download_aos()
    n_steps = int((T(t_final) - T(t_start)) / T(tau))
    if fout:
        fout.write("%s;%d;%d;%d;%g;%g\n" % ("double" if precision == 64 else "float", n, n_steps,
                                              128, total_ms / 1e3, total_ms / 1e3 / max(n_steps, 1)))
        fout.close()
    if fout_1:
        fout_1.close()
    return final

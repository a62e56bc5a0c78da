#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 N-body hot path.

Metric (BASELINE.json): G body-interactions/s, FP64 RK4, N = 1 048 576 Plummer sphere
(configs[3]; it fits one GPU, so N=1 runs the same workload), self pairs counted as the
reference's loop bounds do: interactions = stages * N^2 * steps.

  python bench.py [--gpus N] [--steps K] [--warmup W]            # this framework (CUDA)
  python bench.py --impl reference [--gpus N] [--steps K] ...    # the reference's CPU path

For N > 1 launch under torchrun (one rank per GPU); strong scaling (N^2 work is fixed).
One "step" is one RK4 step (4 force evaluations + 4 fused stage kernels, and for N > 1 one
NCCL all-gather per stage).  `value` is timed with CUDA events on the library's compute
stream (state resident in HBM), max over ranks; `e2e` goes through the public C-ABI calls
with pinned HOST buffers (upload + step + download inside the timed region).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

FLOPS_PER_INTERACTION = 20.0          # BASELINE.json / SURVEY.md 8d
FP64_DATASHEET_TFLOPS = 148 * 64 * 2 * 1.965e9 / 1e12   # 37.2: 64 DFMA lanes/SM/clk @1965 MHz


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--bodies", "--n", dest="n", type=int, default=1 << 20,
                    help="number of bodies (use --bodies under torchrun: its parser claims --n*)")
    ap.add_argument("--precision", type=int, default=64, choices=[64, 32])
    ap.add_argument("--scheme", default="rk4", choices=["rk2", "rk4"])
    ap.add_argument("--dist", default="plummer", choices=["plummer", "uniform"])
    ap.add_argument("--variant", type=int, default=-1)
    ap.add_argument("--unimass", type=int, default=-1, help="-1 auto-detect equal masses, 0 off")
    ap.add_argument("--p2p", type=int, default=0,
                    help="1: peer-pull exchange (TMA reads of the peers' tiles) instead of NCCL all-gathers")
    ap.add_argument("--pair-symmetry", type=int, default=-1,
                    help="-1: the library's automatic choice (pair-symmetric kernel for N >= 2048); "
                         "1: evaluate each pair of distinct bodies once and credit both; "
                         "0: the one-sided kernel (every ordered pair, as the reference does; "
                         "bit-reproducible)")
    ap.add_argument("--sym-variant", type=int, default=-1)
    ap.add_argument("--deterministic", type=int, default=0,
                    help="1: the pair-symmetric kernel with fixed-point row sums (bit-reproducible "
                         "whatever the scheduling); 0 (default): floating-point RED sums")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-energy", action="store_true")
    ap.add_argument("--trace-out", default="",
                    help="write a per-rank operation list (CUDA-event times of one extra, untimed step: "
                         "force launches, stage kernels, memsets, all-gathers, reduce-scatters) to this JSON file")
    ap.add_argument("--no-secondary", action="store_true",
                    help="skip the 2-step general-mass (unimass=0) figure")
    ap.add_argument("--no-parity", action="store_true",
                    help="skip the post-run check of owned RHS rows against the oracle")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    return ap.parse_args()


def make_bodies(args):
    from parallel_comp_labs_b200 import generators
    if args.dist == "plummer":
        # Plummer, a=1, G*M=1, eps = 1e-3 (clamp radius), tau = 1e-3
        return generators.plummer_bodies(args.n, seed=0x5EED0004), 1e-3, 1e-3
    return generators.uniform_bodies(args.n, seed=0x5EED0001), 0.1, 1e-6


# ---------------------------------------------------------------- clocks sampler
class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU while the timed region runs."""

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._th = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksThrottleReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if mask & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            self._stop.wait(0.2)

    def start(self):
        if self.nv:
            self._th = threading.Thread(target=self._run, daemon=True)
            self._th.start()

    def stop(self):
        if self._th:
            self._stop.set()
            self._th.join()
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ---------------------------------------------------------------- CPU baseline (reference)
def cpu_reference_sample(bodies, eps, seconds, steps=1, warmup=0):
    """Times the reference's own CPU implementation of the path on this box's host cores:
    oracle/_ref (the reference's lab_4 sources compiled unmodified) when that build
    travelled here, else the C port.  A BOUNDED sample of the same workload: P thread-ranks
    each evaluate a block of i-rows of one RK stage's RHS against all N bodies through the
    reference's N_body_func(traj, diff, begin, end).  Returns (G int/s per step list, info)."""
    from oracle import Oracle, RefLab4
    n = bodies.shape[0]
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else os.cpu_count()
    cores = max(1, min(cores, n))
    use_ref = RefLab4.available()
    if use_ref:
        ref = RefLab4()
        run = lambda rows: ref.rhs_ranks(bodies, rows, cores, eps=eps)[0]
        kind = "reference"
    else:
        o = Oracle()

        def run(rows):
            t0 = time.perf_counter()
            o.rhs(bodies, 0, rows * cores, eps=eps, threads=cores)
            return time.perf_counter() - t0
        kind = "port"
    rows = 1
    t = run(rows)
    per_row = max(t / rows, 1e-9)
    rows = int(max(1, min(n // cores, seconds / max(steps + warmup, 1) / per_row)))
    rates = []
    for k in range(warmup + steps):
        t = run(rows)
        if k >= warmup:
            rates.append(cores * rows * n / t / 1e9)
    info = {"cores": cores, "kind": kind, "unit": "G interactions/s",
            "sample": f"one RK-stage RHS: rows [0,{cores * rows}) of N={n} against all N bodies, "
                      f"{cores} thread-ranks x {rows} rows through the reference's N_body_func "
                      f"(lab_4 struct_ode_system.hpp:105-128), FP64"}
    return rates, info


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # the CPU arm runs once, on rank 0
    bodies, tau, eps = make_bodies(args)
    rates, info = cpu_reference_sample(bodies, eps, seconds=max(20.0, 4.0 * (args.steps + args.warmup)),
                                       steps=args.steps, warmup=args.warmup)
    value = float(np.mean(rates))
    stages = 4 if args.scheme == "rk4" else 2
    n = args.n
    line = {
        "impl": "reference", "metric": "G body-interactions/s", "value": value,
        "unit": "G interactions/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": stages * n * n / (value * 1e9) * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": workload_config(args, tau, eps, "n/a"),
        "cpu_baseline": dict(info, value=value),
        "e2e": {"value": value, "unit": "G interactions/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "extrapolated_from_rows": True,
        "note": "reference CPU path (lab_4) on host cores, timed on a bounded ROW SAMPLE of one RK "
                "stage's RHS (cpu_baseline.sample); ms_per_step is NOT a measured step: it is the time "
                "a full step of this workload would take at the sampled rate",
    }
    print(json.dumps(line), flush=True)


def workload_config(args, tau, eps, variant):
    shape = "Plummer sphere" if args.dist == "plummer" else "uniform-random cube"
    return {"workload": f"{shape} N={args.n} FP{args.precision} {args.scheme.upper()} "
                        f"tau={tau:g} eps={eps:g} G=6.67e-11 (BASELINE.json configs[3], "
                        f"strong scaling over --gpus)",
            "n_bodies": args.n, "scheme": args.scheme, "tau": tau, "eps": eps,
            "l2": "flushed between timed steps (256 MiB device write); working set 32 B/body",
            "parallelism": (f"i-rows over {args.gpus} GPU(s), peer-pull exchange: remote j-tiles "
                            f"fetched from the peers' HBM by the force kernel's bulk-async copies"
                            if getattr(args, "p2p", 0) and args.gpus > 1 else
                            f"pair-symmetric kernel: the blocks of one global schedule split "
                            f"over {args.gpus} GPU(s); per RK stage an NCCL all-gather of the packed "
                            f"positions (waited for) and a reduce-scatter of the acceleration sums"
                            if getattr(args, "sym_in_use", False) else
                            f"i-rows over {args.gpus} GPU(s), NCCL all-gather of packed positions "
                            f"per RK stage overlapped with the local-tile force launch"),
            "kernel_variant": variant}


# ---------------------------------------------------------------- post-run parity (checker only)
def parity_block(ctx, bodies, eps, precision, rows_per_run=32, runs=4, tau=None):
    """After the timed region: this rank evaluates one RHS of the bench's initial state through
    the library and checks `runs` x `rows_per_run` of ITS OWNED rows against the oracle (the
    reference's begin,end API, struct_ode_system.hpp:53-57: rows [b,e) against all N bodies).
    The oracle is used here as the checker only; nothing it computes is timed or shipped.
    Returns (rows, norm_relative_error, worst_per_row_error)."""
    from oracle import Oracle
    o = Oracle()
    n = bodies.shape[0]
    dtype = np.float64 if precision == 64 else np.float32
    ctx.upload_aos(np.ascontiguousarray(bodies, dtype=dtype))
    got = ctx.rhs().astype(np.float64)
    b, e = ctx.owned_range()
    own = e - b
    rows_per_run = max(0, min(rows_per_run, own))
    # (a rank that owns no rows checks nothing but still takes part in the collectives below)
    starts = (sorted({b + int(k * (own - rows_per_run) / max(runs - 1, 1)) for k in range(runs)})
              if own > 0 else [])
    worst_norm = worst_row = 0.0
    ref64 = np.ascontiguousarray(bodies.astype(dtype), dtype=np.float64)  # the values the GPU holds
    for s0 in starts:
        want = o.rhs(ref64, s0, s0 + rows_per_run, eps=eps, threads=2)[s0:s0 + rows_per_run, 4:7]
        g = got[s0:s0 + rows_per_run]
        worst_norm = max(worst_norm, float(np.abs(g - want).max() / np.abs(want).max()))
        worst_row = max(worst_row, float((np.abs(g - want).max(axis=1)
                                          / np.maximum(np.abs(want).max(axis=1), 1e-300)).max()))
    PARITY_EXTRA.clear()
    if tau is not None:
        # One 1-stage step (Euler, through nbody_step_tableau: the fused stage kernel, the end-of-step
        # exchange of Y on several GPUs) and a second RHS on the stepped, gathered state:
        #   y1 = y0 + tau f(y0): sampled owned rows against y0 and the oracle rows computed above;
        #   f(y1): the same rows against the oracle evaluated on the DOWNLOADED y1 (all N bodies).
        ctx.step_tableau([], [1.0], 1.0, tau, 1, sync=True)
        y1 = ctx.download_aos().astype(np.float64)      # collective: every rank calls parity_block
        got2 = ctx.rhs().astype(np.float64)
        st_err = dv_err = r2_err = 0.0
        for s0 in starts:
            sl = slice(s0, s0 + rows_per_run)
            a0 = o.rhs(ref64, s0, s0 + rows_per_run, eps=eps, threads=2)[sl, 4:7]
            want_r = ref64[sl, 1:4] + tau * ref64[sl, 4:7]
            want_v = ref64[sl, 4:7] + tau * a0
            st_err = max(st_err, float(np.abs(y1[sl, 1:4] - want_r).max() / np.abs(want_r).max()),
                         float(np.abs(y1[sl, 4:7] - want_v).max() / np.abs(want_v).max()))
            dv_err = max(dv_err, float(np.abs((y1[sl, 4:7] - ref64[sl, 4:7]) - tau * a0).max()
                                       / np.abs(tau * a0).max()))
            want2 = o.rhs(y1, s0, s0 + rows_per_run, eps=eps, threads=2)[sl, 4:7]
            r2_err = max(r2_err, float(np.abs(got2[sl] - want2).max() / np.abs(want2).max()))
        PARITY_EXTRA.update(state_rel_err=st_err, dv_rel_err=dv_err, rhs2_rel_err=r2_err)
    return len(starts) * rows_per_run, worst_norm, worst_row


PARITY_EXTRA = {}     # filled by parity_block(tau=...): state after one Euler step, second RHS
PARITY_TOL = {64: (1e-12, 1e-10), 32: (1e-4, 2e-3)}     # (norm-relative, per-row) -- tests/ hold the same


# ---------------------------------------------------------------- this framework
def run_b200(args):
    import torch
    import torch.distributed as dist
    from parallel_comp_labs_b200 import capi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("NCCL_DEBUG", "WARN")   # keep stdout to the one JSON line
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus N > 1 must be launched with torch.distributed.run "
                             "(--nproc-per-node N)")
        args.gpus = world
    torch.cuda.set_device(local)
    uid = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        box = [capi.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        uid = box[0]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    bodies, tau, eps = make_bodies(args)
    n = args.n
    scheme = capi.NBODY_RK4 if args.scheme == "rk4" else capi.NBODY_RK2
    stages = 4 if args.scheme == "rk4" else 2
    dtype = np.float64 if args.precision == 64 else np.float32

    peak = capi.measure_fma_peak(args.precision, 300.0, device=local)   # roofline denominator
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    ctx = capi.Context(n, args.precision, local, rank, world, uid)
    if args.p2p and world > 1:
        blobs = [None] * world
        dist.all_gather_object(blobs, ctx.p2p_export())
        ctx.p2p_connect(blobs)
    ctx.set_params(6.67e-11, eps)
    ctx.set_option("variant", args.variant)
    ctx.set_option("unimass", args.unimass)
    ctx.set_option("pair_symmetry", args.pair_symmetry)
    ctx.set_option("sym_variant", args.sym_variant)
    ctx.set_option("deterministic", args.deterministic)
    host_in = capi.PinnedArray((n, 7), dtype)
    host_out = capi.PinnedArray((n, 7), dtype)
    host_in.array[:] = bodies.astype(dtype)
    ctx.upload_aos(host_in.array)
    e0 = None if args.no_energy else sum(ctx.energy())

    # ---- device-resident timed region: W warm-up steps, then K timed steps ------------------
    for _ in range(args.warmup):
        ctx.step(scheme, tau, 1)
    ctx.sync()
    ctx.set_option("time_kernels", 1)
    launches0 = ctx.timing().total_launches
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    step_ms, force_ms, force_launches, pair_evals = [], 0.0, 0, 0
    for _ in range(args.steps):
        flush.zero_()                      # L2 flush between timed iterations (untimed)
        torch.cuda.synchronize()
        barrier()
        ctx.step(scheme, tau, 1)
        ctx.sync()
        tm = ctx.timing()                  # CUDA events on the library's compute stream
        step_ms.append(max_over_ranks(tm.last_call_ms))
        force_ms += tm.force_ms
        force_launches += tm.force_launches
        pair_evals += tm.pair_evaluations
    barrier()
    clocks = sampler.stop()
    launches = ctx.timing().total_launches - launches0
    ctx.set_option("time_kernels", 0)
    total_ms = float(np.sum(step_ms))
    value = stages * float(n) * float(n) * args.steps / (total_ms * 1e-3) / 1e9
    e1 = None if args.no_energy else sum(ctx.energy())

    # roofline of the dominant kernel (force): algorithmic flops per launch / mean duration
    b, e = ctx.owned_range()
    inter_per_launch = (e - b) * float(n) * stages * args.steps / max(force_launches, 1)
    mean_force_ms = force_ms / max(force_launches, 1)
    achieved = FLOPS_PER_INTERACTION * inter_per_launch / (mean_force_ms * 1e-3) / 1e12
    force_share = max_over_ranks(force_ms) / total_ms

    # ---- end-to-end through the public API with pinned host buffers --------------------------
    # One rank: the whole state goes up and comes back (nbody_upload_aos / nbody_download_aos).
    # Several ranks: each rank moves only the rows it owns through PCIe (nbody_upload_aos_owned /
    # nbody_download_aos_owned, as an MPI rank of lab_4 owns rows [offsets[id], +counts[id]));
    # the full position array is assembled by an all-gather over NVLink.  Bytes per step are the
    # same in total; they are the sum over ranks of what each rank copies.
    e2e_ms = []
    ob, oe = ctx.owned_range()
    own_in = host_in.array[ob:oe]
    own_out = host_out.array[ob:oe]
    for k in range(1 + args.steps):        # first pass untimed
        barrier()
        t0 = time.perf_counter()
        if world == 1 or args.p2p:
            ctx.upload_aos(host_in.array)      # H2D of the step's input state
            ctx.step(scheme, tau, 1)
            ctx.download_aos(host_out.array)   # D2H of the step's result
        else:
            ctx.upload_aos_owned(own_in)
            ctx.step(scheme, tau, 1)
            ctx.download_aos_owned(own_out)
        dt = (time.perf_counter() - t0) * 1e3
        if k:
            e2e_ms.append(max_over_ranks(dt))
    e2e_value = stages * float(n) * float(n) * len(e2e_ms) / (sum(e2e_ms) * 1e-3) / 1e9
    bytes_io = n * 7 * np.dtype(dtype).itemsize * (world if (world > 1 and args.p2p) else 1)
    # ---- per-rank operation list of one extra step (untimed; the library's own event pairs) ----
    if args.trace_out:
        ctx.set_option("trace", 1)
        barrier()
        ctx.step(scheme, tau, 1)
        ctx.sync()
        mine = {"rank": rank, "step_ms": ctx.timing().last_call_ms, "ops": ctx.trace()}
        ctx.set_option("trace", 0)
        if world > 1:
            allr = [None] * world
            dist.all_gather_object(allr, mine)
        else:
            allr = [mine]
        if rank == 0:
            with open(args.trace_out, "w") as f:
                json.dump({"workload": workload_config(args, tau, eps, None)["workload"], "n_gpus": world,
                           "ranks": allr}, f, indent=1)
    # ---- parity of THIS run's configuration (every rank checks rows it owns; not timed) -------
    parity = None
    if not args.no_parity:
        prow, pnorm, pper = parity_block(ctx, bodies, eps, args.precision, tau=tau)
        extra = {k: max_over_ranks(v) for k, v in sorted(PARITY_EXTRA.items())}
        tol_norm, tol_row = PARITY_TOL[args.precision]
        tol_state = 1e-10 if args.precision == 64 else 1e-4      # north_star's trajectory tolerance
        parity = {"rows": int(sum_over_ranks(float(prow))), "rows_per_rank": int(prow),
                  "rhs_rel_err": max_over_ranks(pnorm), "per_row_max": max_over_ranks(pper),
                  "tolerance": {"rhs_rel_err": tol_norm, "per_row": tol_row},
                  "checker": "oracle/nbody_oracle.c (bit-exact restatement of lab_4 N_body_func), "
                             "rows [b,e) of each rank's owned block against all N bodies"}
        parity["ok"] = bool(parity["rhs_rel_err"] < tol_norm and parity["per_row_max"] < tol_row)
        if extra:
            # (dv_rel_err is reported, not gated: v1 - v0 cancels, its floor is eps_mach |v| / |tau a|)
            parity.update(extra)
            parity["after_one_euler_step"] = ("state_rel_err: positions / velocities of sampled owned rows "
                                              "after y1 = y0 + tau f(y0) (nbody_step_tableau, 1 stage); "
                                              "dv_rel_err: their velocity increments against tau x oracle rows; "
                                              "rhs2_rel_err: nbody_rhs on y1 against the oracle on the downloaded y1")
            parity["tolerance"].update(state_rel_err=tol_state, rhs2_rel_err=tol_norm)
            parity["ok"] = bool(parity["ok"] and extra["state_rel_err"] < tol_state
                                and extra["rhs2_rel_err"] < tol_norm)
    # ---- secondary figure: the same workload with the uniform-mass specialisation OFF ----------
    # (the Plummer bodies have equal masses, which the library detects and factors out of the
    # pair sum; inputs with distinct masses -- the reference's own rand_bodies.txt -- take the
    # general-mass kernels: one more FP64 instruction per pair / two more per pair evaluation)
    secondary = None
    if not args.no_secondary and tm.uniform_mass and args.unimass != 0:
        uid2 = None
        if world > 1:
            box = [capi.nccl_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(box, src=0)
            uid2 = box[0]
        with capi.Context(n, args.precision, local, rank, world, uid2) as c2:
            c2.set_params(6.67e-11, eps)
            c2.set_option("unimass", 0)
            c2.set_option("pair_symmetry", args.pair_symmetry)
            c2.upload_aos(host_in.array)
            c2.step(scheme, tau, 1)
            c2.sync()
            ms2 = []
            for _ in range(2):
                barrier()
                c2.step(scheme, tau, 1)
                c2.sync()
                ms2.append(max_over_ranks(c2.timing().last_call_ms))
            t2 = c2.timing()
            secondary = {"general_masses": {
                "value": stages * float(n) * float(n) * 2 / (sum(ms2) * 1e-3) / 1e9,
                "unit": "G interactions/s", "steps": 2, "warmup": 1,
                "uniform_mass_factored_out": bool(t2.uniform_mass),
                "pair_symmetry": bool(t2.pair_symmetry),
                "note": "same bodies, option unimass=0: the kernels any input with distinct masses gets"}}
    # ---- secondary figure: the bit-reproducible mode of the pair-symmetric kernel (one GPU) -----
    # option "deterministic": fixed-point row sums.  Two RHS evaluations must agree bit for bit.
    if (not args.no_secondary and world == 1 and tm.pair_symmetry and not args.deterministic):
        try:
            with capi.Context(n, args.precision, local, rank, world, None) as c3:
                c3.set_params(6.67e-11, eps)
                c3.set_option("unimass", args.unimass)
                c3.set_option("pair_symmetry", args.pair_symmetry)
                c3.set_option("deterministic", 1)
                c3.upload_aos(host_in.array)
                r1 = c3.rhs()
                r2 = c3.rhs()
                c3.step(scheme, tau, 1)
                c3.sync()
                ms3 = []
                for _ in range(2):
                    c3.step(scheme, tau, 1)
                    c3.sync()
                    ms3.append(c3.timing().last_call_ms)
                t3 = c3.timing()
                det = {"value": stages * float(n) * float(n) * 2 / (sum(ms3) * 1e-3) / 1e9,
                       "unit": "G interactions/s", "steps": 2, "warmup": 1,
                       "deterministic": bool(t3.deterministic), "sym_variant": int(t3.sym_variant),
                       "two_rhs_bitwise_equal": bool(np.array_equal(r1, r2)),
                       "note": "option deterministic=1: every unit's share of a row is added as an exact "
                               "fixed-point integer (64-bit integer RED) instead of a floating-point RED"}
        except capi.NBodyError as ex:
            det = {"error": str(ex)}
        secondary = dict(secondary or {}, deterministic=det)
    # ---- the other BASELINE.json configs, short, after the timed region (secondary.configs) -----
    # The headline line above is configs[3] FP64.  So that the driver's BENCH / SCALE records also
    # hold the others, each gets a few device-timed steps here, with its own parity check against
    # the oracle (owned rows of one RHS) -- same timing rules (warm-up >= 3, L2 flushed between
    # steps, CUDA events on the library's stream, max over ranks), a fresh context each.
    def other_config(precision, n_bodies, dist_name, steps, warmup, with_cpu=False):
        sub = argparse.Namespace(**vars(args))
        sub.n, sub.dist, sub.precision = n_bodies, dist_name, precision
        sub.variant, sub.sym_variant, sub.pair_symmetry, sub.unimass, sub.deterministic = -1, -1, -1, -1, 0
        bod, tau_s, eps_s = make_bodies(sub)
        uid_s = None
        if world > 1:
            box = [capi.nccl_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(box, src=0)
            uid_s = box[0]
        dt_s = np.float64 if precision == 64 else np.float32
        try:
            with capi.Context(n_bodies, precision, local, rank, world, uid_s) as c:
                c.set_params(6.67e-11, eps_s)
                c.upload_aos(np.ascontiguousarray(bod, dtype=dt_s))
                for _ in range(warmup):
                    c.step(scheme, tau_s, 1)
                c.sync()
                ms = []
                for _ in range(steps):
                    flush.zero_()
                    torch.cuda.synchronize()
                    barrier()
                    c.step(scheme, tau_s, 1)
                    c.sync()
                    ms.append(max_over_ranks(c.timing().last_call_ms))
                t = c.timing()
                prow, pnorm, pper = parity_block(c, bod, eps_s, precision, rows_per_run=32, runs=2)
            tol_norm, tol_row = PARITY_TOL[precision]
            pn, pp = max_over_ranks(pnorm), max_over_ranks(pper)
            val = stages * float(n_bodies) ** 2 * steps / (sum(ms) * 1e-3) / 1e9
            shape = "Plummer sphere" if dist_name == "plummer" else "uniform-random cube"
            out = {"workload": f"{shape} N={n_bodies} FP{precision} {args.scheme.upper()} tau={tau_s:g} "
                               f"eps={eps_s:g} G=6.67e-11, L2 flushed between timed steps",
                   "value": val, "unit": "G interactions/s", "n_gpus": world, "steps": steps,
                   "warmup": warmup, "ms_per_step": float(np.mean(ms)),
                   "dtype": "f64" if precision == 64 else "f32",
                   "pair_symmetry": bool(t.pair_symmetry), "sym_variant": int(t.sym_variant),
                   "uniform_mass_factored_out": bool(t.uniform_mass),
                   "parity": {"rows": int(sum_over_ranks(float(prow))), "rhs_rel_err": pn,
                              "per_row_max": pp, "ok": bool(pn < tol_norm and pp < tol_row)}}
            if with_cpu and rank == 0:
                # small enough for the reference's CPU code to evaluate a WHOLE RHS, not a row sample
                rates, info = cpu_reference_sample(bod.astype(np.float64), eps_s, seconds=1e9, steps=2, warmup=1)
                out["cpu_baseline"] = dict(info, value=float(np.mean(rates)),
                                           whole_rhs=bool(f"rows [0,{n_bodies})" in info["sample"]))
            return out
        except capi.NBodyError as ex:
            return {"error": str(ex)}

    if not args.no_secondary and args.n == (1 << 20) and args.dist == "plummer" and args.precision == 64 \
            and args.scheme == "rk4" and not args.p2p:
        cfgs = {"plummer1m_fp32": dict(other_config(32, 1 << 20, "plummer", 3, 3),
                                       baseline_config="configs[3], FP32 half"),
                "plummer256k_fp64": dict(other_config(64, 1 << 18, "plummer", 5, 3),
                                         baseline_config="configs[2] (strong scaling over --gpus)")}
        if world == 1:
            cfgs["uniform20k_fp64"] = dict(other_config(64, 20000, "uniform", 10, 3,
                                                        with_cpu=not args.no_cpu_baseline),
                                           baseline_config="configs[1] (1 GPU vs the lab_4 CPU path)")
        secondary = dict(secondary or {}, configs=cfgs)
    total_launches = int(sum_over_ranks(float(launches)))
    sum_pairs = sum_over_ranks(float(pair_evals))

    args.sym_in_use = bool(tm.pair_symmetry)
    line = None
    if rank == 0:
        variant_name = {"variant": int(tm.variant), "uniform_mass_factored_out": bool(tm.uniform_mass),
                        "fp64_instr_per_pair": ((15 if tm.uniform_mass else 16)
                                                - (1 if tm.variant in (5, 6) else 0))
                        if args.precision == 64 else None}
        if tm.pair_symmetry:
            # every pair evaluation serves two interactions: 18 (20 with general masses) FP64
            # instructions per evaluation; `value` and roofline.achieved count DELIVERED
            # interactions (the metric's definition), roofline.evaluated_* what was computed
            variant_name.update({"pair_symmetry": True, "sym_variant": int(tm.sym_variant),
                                 "deterministic_row_sums": bool(tm.deterministic),
                                 "fp64_instr_per_pair": None,
                                 "fp64_instr_per_pair_evaluation":
                                     (18 if tm.uniform_mass else 20) if args.precision == 64 else None})
        archived = archived_ncu(args.n, world, args.precision, bool(tm.pair_symmetry),
                                bool(tm.uniform_mass),
                                int(tm.sym_variant) if tm.pair_symmetry else int(tm.variant))
        line = {
            "metric": "G body-interactions/s", "value": value, "unit": "G interactions/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64" if args.precision == 64 else "f32",
            "data": "synthetic", "config": workload_config(args, tau, eps, variant_name),
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "G interactions/s",
                    "h2d_bytes_per_step": int(bytes_io), "d2h_bytes_per_step": int(bytes_io),
                    "ms_per_step": float(np.mean(e2e_ms))},
            "gpu_launches": total_launches,
            "roofline": {
                "bound": "fp64_fma_pipe" if args.precision == 64 else "fp32_fma_pipe",
                "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                "traffic": archived["traffic"],
                "traffic_source": archived["traffic_source"],
                "algorithmic_bytes_per_launch": int(n * 4 * np.dtype(dtype).itemsize
                                                    + (e - b) * 3 * np.dtype(dtype).itemsize),
                "pipe_util_ncu_archived": archived["pipe_util"],
                "kernel": "nbody::force_sym_kernel" if tm.pair_symmetry else "nbody::force_kernel",
                "mean_launch_ms": mean_force_ms,
                "evaluated_pairs_per_delivered_interaction":
                    float(sum_pairs) / (stages * float(n) * float(n) * args.steps),
                "share_of_step": force_share,
                "flops_per_interaction": FLOPS_PER_INTERACTION,
                "peak_source": "measured live: nbody_measure_fma_peak (register-resident DFMA "
                               "chains, 300 ms); MEASURED_PEAKS.json has no FP64 vector figure",
                "frac_of_datasheet_fp64": (achieved / FP64_DATASHEET_TFLOPS
                                           if args.precision == 64 else None),
                "note": "compute-bound path (6e5 flop/B): neither HBM nor tensor cores bound it",
            },
            "pct_of_fp_peak_whole_job": 100.0 * FLOPS_PER_INTERACTION * value * 1e9 / 1e12
                                        / (peak * world),
            "energy_drift": (abs(e1 - e0) / abs(e0)) if e0 else None,
            "parity": parity,
            "secondary": secondary,
        }
    ctx.close()
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        rates, info = cpu_reference_sample(bodies.astype(np.float64), eps, args.cpu_seconds)
        line["cpu_baseline"] = dict(info, value=float(np.mean(rates)))
    elif rank == 0:
        line["cpu_baseline"] = None
    if rank == 0:
        print(json.dumps(line), flush=True)
    host_in.free()
    host_out.free()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if parity is not None and not parity["ok"]:
        raise SystemExit(f"PARITY FAILED: {parity}")


# ARCHIVED profiler readings of the dominant kernel, from the committed `ncu --set full` captures
# under profiles/ -- ncu cannot run inside a timed bench, so these are NOT measurements of this run.
# A line carries them only when the run is the captured configuration: same n, one GPU, same
# precision, same kernel and kernel variant, same mass mode; anything else gets null.
#   traffic   = dram__bytes_read.sum + dram__bytes_write.sum per force-kernel launch
#   pipe_util = sm__inst_executed_pipe_fp64 / _fma .avg.pct_of_peak_sustained_active
# key: (n, gpus, precision, pair-symmetric kernel?, equal masses?, variant index)
ARCHIVED_NCU = {
    (1 << 20, 1, 64, False, True, 0): {"traffic": 83091968 + 963913216, "pipe_util": 0.8974,
                                       "source": "profiles/r01_force_kernel_1M_ncu_full.csv"},
    (1 << 20, 1, 64, True, True, 8): {"traffic": 58879744 + 550656, "pipe_util": 0.8377,
                                      "source": "profiles/r02_ncu_k1s_fp64_equal_1M.csv"},
    (1 << 18, 1, 64, True, True, 9): {"traffic": 14749184 + 11520, "pipe_util": None,
                                      "source": "profiles/r02_ncu_k1s_fp64_equal_262144.csv"},
    (1 << 18, 1, 64, True, False, 5): {"traffic": None, "pipe_util": 0.8248,
                                       "source": "profiles/r02_ncu_k1s_fp64_general_262144.csv"},
    (1 << 18, 1, 32, True, True, 1): {"traffic": 7392768 + 0, "pipe_util": 0.6923,
                                      "source": "profiles/r02_ncu_k1s_fp32_262144.csv",
                                      "note": "issue-bound: 83 % of issue slots"},
    (20000, 1, 64, False, False, 0): {"traffic": None, "pipe_util": 0.8195,
                                      "source": "profiles/r02_ncu_k1_fp64_general_20000.csv"},
    (131072, 1, 32, False, True, 0): {"traffic": None, "pipe_util": 0.7031,
                                      "source": "profiles/r02_ncu_k1_fp32_131072.csv"},
}


def archived_ncu(n, gpus, precision, sym, unimass, variant):
    hit = ARCHIVED_NCU.get((n, gpus, precision, sym, unimass, variant))
    if hit is None:
        return {"traffic": None, "pipe_util": None,
                "traffic_source": "no committed ncu capture of this configuration (n, GPUs, precision, "
                                  "kernel variant, mass mode); ncu cannot run inside a timed bench"}
    pu = None
    if hit["pipe_util"] is not None:
        pu = {"value": hit["pipe_util"], "source": hit["source"], "captured_variant": variant,
              "captured_n": n, "live": False}
        if "note" in hit:
            pu["note"] = hit["note"]
    return {"traffic": hit["traffic"], "pipe_util": pu,
            "traffic_source": "archived, not live: " + hit["source"]}


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)

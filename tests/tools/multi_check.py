"""Multi-GPU parity check, run under torchrun (one rank per GPU):
   python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
       --master-port 29511 tests/tools/multi_check.py [n]
Every rank owns a block of i-rows; the packed positions are NCCL-all-gathered per RK stage.
Checks against the oracle (RHS rows, RK2 + RK4 states) and against a single-GPU run."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
from parallel_comp_labs_b200 import capi, generators, tableaux
from oracle import Oracle

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
def fresh_uid():  # one NCCL unique id per communicator (= per context)
    box = [capi.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    return box[0]
n = int(sys.argv[1]) if len(sys.argv) > 1 else 5000
P2P = os.environ.get("NBODY_P2P", "0") == "1"
# NBODY_PAIR_SYM: "1" pair-symmetric kernel + reduce-scatter, "0" one-sided kernel (all-gather
# overlapped with the local-tile launch), unset / "auto" = the library's choice
SYM_OPT = {"0": 0, "1": 1}.get(os.environ.get("NBODY_PAIR_SYM", "auto"), -1)
# NBODY_DETERMINISTIC=1: option "deterministic" (fixed-point row sums, integer reduce-scatter): the
# runs must then agree BIT FOR BIT where the floating-point sums agree to rounding
DET = os.environ.get("NBODY_DETERMINISTIC", "0") == "1"
PAIR_SYM = SYM_OPT != 0 and not DET  # tolerance class of the owned-rows comparison (atomics: rounding only)
o = Oracle()
ok = True
for (name, bodies, eps, tau) in (("uniform", generators.uniform_bodies(n, 31), 1e-6, 0.1),
                                 ("plummer", generators.plummer_bodies(n + 37, 32), 1e-3, 1e-3)):
    nn = bodies.shape[0]
    for prec, tol in ((64, 1e-12), (32, 1e-4)):
        with capi.Context(nn, prec, local, rank, world, fresh_uid()) as ctx:
            if P2P:  # peer-pull exchange instead of NCCL all-gathers
                blobs = [None] * world
                dist.all_gather_object(blobs, ctx.p2p_export())
                ctx.p2p_connect(blobs)
            ctx.set_params(6.67e-11, eps)
            ctx.set_option("pair_symmetry", SYM_OPT)
            ctx.set_option("deterministic", 1 if DET else 0)
            ctx.upload_aos(bodies)
            b, e = ctx.owned_range()
            a = ctx.rhs().astype(np.float64)
            a_again = ctx.rhs().astype(np.float64) if DET else None
            want = o.rhs(bodies, b, e, eps=eps, threads=4)[b:e, 4:7]
            full = o.rhs(bodies, eps=eps, threads=4)[:, 4:7] if rank == 0 and nn <= 20000 else None
            err = np.abs(a[b:e] - want).max() / max(np.abs(want).max(), 1e-300) if e > b else 0.0
            assert np.all(np.isnan(a[:b])) and np.all(np.isnan(a[e:]))
            res = {"rhs": err, "sym": float(ctx.timing().pair_symmetry)}
            if DET:
                assert ctx.timing().deterministic == 1
                res["det_bits_differ"] = 0.0 if np.array_equal(a[b:e], a_again[b:e]) else 1.0
            for scheme, fn in ((2, o.rk2), (4, o.rk4)):
                ctx.upload_aos(bodies)
                ctx.step(scheme, tau, 3)
                snap = ctx.snapshot()
                fin = ctx.download_aos().astype(np.float64)
                k, u = ctx.energy()
                w, _ = fn(bodies, tau, 0., 1.0, eps=eps, max_steps=3, rows=False, threads=4)
                res[f"rk{scheme}_r"] = np.abs(fin[:, 1:4] - w[:, 1:4]).max() / np.abs(w[:, 1:4]).max()
                dv = w[:, 4:7] - bodies[:, 4:7]
                res[f"rk{scheme}_dv"] = np.abs(fin[:, 4:7] - bodies[:, 4:7] - dv).max() / np.abs(dv).max()
                # cancellation floor of v(t)-v(0): 4 roundings per step of size eps_mach*|v|
                dv_floor = 12 * (2.2e-16 if prec == 64 else 1.2e-7) * np.abs(w[:, 4:7]).max() / np.abs(dv).max()
                res[f"rk{scheme}_snap"] = float(np.abs(snap - fin[:, 1:4]).max())
                ko, uo = o.energy(w, eps=eps, threads=4)
                res[f"rk{scheme}_E"] = abs((k + u) - (ko + uo)) / abs(ko + uo)
                if not P2P:
                    # the same run with every rank moving only the rows it owns through PCIe
                    ctx.upload_aos_owned(bodies[b:e].astype(np.float64 if prec == 64 else np.float32))
                    ctx.step(scheme, tau, 3)
                    own = ctx.download_aos_owned().astype(np.float64)
                    res[f"rk{scheme}_owned"] = (float(np.abs(own - fin[b:e]).max()
                                                      / np.abs(fin[:, 1:]).max()) if e > b else 0.0)
            if prec == 64 and not P2P:
                # K2g on several ranks: one embedded DOPRI5 step (error norms all-reduced with MAX),
                # reject it (undo), take a smaller one -- against the oracle's generic integrator
                tb = tableaux.TABLEAUX["dopri5"]
                ctx.upload_aos(bodies)
                nrm = ctx.step_embedded(tb["a"], tb["b"], tb["b_err"], 2 * tau)
                w1, werr = o.rk_general(bodies, tb["a"], tb["b"], 2 * tau, 1, b_err=tb["b_err"], eps=eps, threads=4)
                res["emb_norms"] = float(np.abs(nrm / werr - 1).max())   # each of the 4 norms, relative
                ctx.step_undo()
                ctx.step_embedded(tb["a"], tb["b"], tb["b_err"], tau)
                fin = ctx.download_aos().astype(np.float64)
                w2, _ = o.rk_general(bodies, tb["a"], tb["b"], tau, 1, eps=eps, threads=4)
                res["emb_r"] = float(np.abs(fin[:, 1:4] - w2[:, 1:4]).max() / np.abs(w2[:, 1:4]).max())
            lim = {"rhs": tol, "_r": tol, "_dv": max(tol * 10, dv_floor), "emb_norms": 1e-4, "_snap": 0.0 if prec == 64 else 1e-6,
                   "_E": max(tol * 10, 1e-9), "_owned": tol if PAIR_SYM else 0.0, "det_bits_differ": 0.0}
            good = all(v <= next(l for suf, l in lim.items() if k_.endswith(suf))
                       for k_, v in res.items() if k_ != "sym")
            ok &= good
            print(f"[rank {rank}/{world}] {name} n={nn} fp{prec} rows[{b},{e}) "
                  + " ".join(f"{k_}={v:.2e}" for k_, v in res.items()) + (" OK" if good else " FAIL"),
                  flush=True)
flag = torch.tensor([0 if ok else 1], device="cuda")
dist.all_reduce(flag)
dist.barrier()
dist.destroy_process_group()
if flag.item():
    print("MULTI-GPU CHECK FAILED"); sys.exit(1)
if rank == 0:
    print("MULTI-GPU CHECK PASSED" + (" (peer-pull exchange)" if P2P else " (NCCL all-gather)")
          + {0: " (one-sided kernel)", 1: " (pair-symmetric kernel)", -1: " (automatic kernel choice)"}[SYM_OPT]
          + (" (deterministic: fixed-point row sums)" if DET else ""))

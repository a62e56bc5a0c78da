"""GPU exploration script (run under gpurun): peak microbenchmarks, parity spot checks,
force-kernel variant sweep.  Writes gpurun_out/explore.json."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from parallel_comp_labs_b200 import capi, generators
from oracle import Oracle

out = {}
o = Oracle()
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)

def dump():
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "explore.json"), "w"), indent=1)

print(capi.lib().nbody_version().decode())
for prec in (64, 32):
    vals = [capi.measure_fma_peak(prec, 300.0) for _ in range(3)]
    out[f"fma_peak_fp{prec}_tflops"] = vals
    print("fma peak", prec, vals, flush=True)
dump()

# parity spot checks
b4 = np.array([[8810324116.227, 1, 0, 0, 0, .9, 0], [8810324116.227, 0, 1, 0, -.9, 0, 0],
               [8810324116.227, -1, 0, 0, 0, -.9, 0], [8810324116.227, 0, -1, 0, .9, 0, 0]])
par = {}
for name, bodies in (("4body", b4), ("uniform3000", generators.uniform_bodies(3000, 7)),
                     ("plummer5000", generators.plummer_bodies(5000, 11))):
    eps = 1e-3 if name.startswith("plummer") else 1e-6
    want = o.rhs(bodies, eps=eps, threads=8)[:, 4:7]
    for prec in (64, 32):
        nv = 7 if prec == 64 else 6
        for var in range(nv):
            with capi.Context(bodies.shape[0], prec) as ctx:
                ctx.set_params(6.67e-11, eps)
                ctx.set_option("variant", var)
                ctx.upload_aos(bodies)
                got = ctx.rhs().astype(np.float64)
            err = np.abs(got - want).max() / np.abs(want).max()
            par[f"{name}/fp{prec}/v{var}"] = err
            print("rhs parity", name, prec, var, err, flush=True)
    for scheme, fn in ((2, o.rk2), (4, o.rk4)):
        tau = 0.1 if not name.startswith("plummer") else 1e-3
        w, _ = fn(bodies, tau, 0., 100., eps=eps, max_steps=5, rows=False, threads=8)
        with capi.Context(bodies.shape[0], 64) as ctx:
            ctx.set_params(6.67e-11, eps)
            ctx.upload_aos(bodies)
            ctx.step(scheme, tau, 5, sync=True)
            g = ctx.download_aos()
            snap = ctx.snapshot()
            en = ctx.energy()
        er = np.abs(g[:, 1:4] - w[:, 1:4]).max() / np.abs(w[:, 1:4]).max()
        dvw, dvg = w[:, 4:7] - bodies[:, 4:7], g[:, 4:7] - bodies[:, 4:7]
        ev = np.abs(dvg - dvw).max() / np.abs(dvw).max()
        eo = o.energy(g, eps=eps, threads=8)
        par[f"{name}/rk{scheme}"] = [er, ev, float(np.abs(snap - g[:, 1:4]).max()),
                                     abs(en[0] - eo[0]) / abs(eo[0]), abs(en[1] - eo[1]) / abs(eo[1])]
        print("rk parity", name, scheme, par[f"{name}/rk{scheme}"], flush=True)
out["parity"] = par
dump()

# variant sweep
sweep = {}
for n in (131072, 20000):
    bodies = generators.plummer_bodies(n, 3)
    for prec in (64, 32):
        nv = 7 if prec == 64 else 6
        for var in range(nv):
            with capi.Context(n, prec) as ctx:
                ctx.set_params(6.67e-11, 1e-3)
                ctx.set_option("variant", var)
                ctx.upload_aos(bodies)
                ts = []
                for rep in range(4):
                    ctx.rhs()
                    ts.append(ctx.timing().last_call_ms)
            rate = n * n / (min(ts) * 1e-3) / 1e9
            sweep[f"n{n}/fp{prec}/v{var}"] = {"ms": ts, "Gint_s": rate}
            print("sweep", n, prec, var, [round(t, 3) for t in ts], round(rate, 1), flush=True)
out["sweep"] = sweep
dump()

# big run: N=1M one RHS + one RK4 step for the default variant
n = 1 << 20
bodies = generators.plummer_bodies(n, 0x5EED0004)
for prec in (64, 32):
    with capi.Context(n, prec) as ctx:
        ctx.set_params(6.67e-11, 1e-3)
        ctx.upload_aos(bodies)
        ctx.rhs(); t1 = ctx.timing().last_call_ms
        ctx.rhs(); t2 = ctx.timing().last_call_ms
        ctx.step(4, 1e-3, 1, sync=True); t3 = ctx.timing().last_call_ms
    out[f"n1M/fp{prec}"] = {"rhs_ms": [t1, t2], "rk4_step_ms": t3,
                            "Gint_s_rhs": n * n / (t2 * 1e-3) / 1e9,
                            "Gint_s_step": 4 * n * n / (t3 * 1e-3) / 1e9}
    print("1M", prec, out[f"n1M/fp{prec}"], flush=True)
    dump()

"""Force-kernel variant sweep on one GPU: RHS parity vs oracle (N=3000) + RHS timing at
several N.  usage: sweep.py [prec] [n ...]   (writes gpurun_out/sweep.json)"""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from parallel_comp_labs_b200 import capi, generators
from oracle import Oracle
precs = [int(sys.argv[1])] if len(sys.argv) > 1 and sys.argv[1] != "all" else [64, 32]
ns = [int(a) for a in sys.argv[2:]] or [131072]
UNIMASS = int(os.environ.get("SWEEP_UNIMASS", "-1"))
o = Oracle()
small = generators.plummer_bodies(3000, 5)
small[7, 1:4] = small[8, 1:4] + 1e-5   # a pair inside the clamp radius
want = o.rhs(small, eps=1e-3, threads=8)[:, 4:7]
res = {}
for prec in precs:
    var = 0
    while True:
        try:
            with capi.Context(3000, prec) as ctx:
                ctx.set_params(6.67e-11, 1e-3)
                ctx.set_option("variant", var)
                ctx.set_option("unimass", UNIMASS)
                ctx.upload_aos(small)
                got = ctx.rhs().astype(np.float64)
        except capi.NBodyError as e:
            if "variant out of range" in str(e): break
            raise
        err = float(np.abs(got - want).max() / np.abs(want).max())
        line = {"rhs_rel_err": err}
        for n in ns:
            bodies = generators.plummer_bodies(n, 3)
            with capi.Context(n, prec) as ctx:
                ctx.set_params(6.67e-11, 1e-3)
                ctx.set_option("variant", var)
                ctx.set_option("unimass", UNIMASS)
                ctx.upload_aos(bodies)
                ts = []
                for rep in range(3 if n >= 500000 else 5):
                    ctx.rhs()
                    ts.append(ctx.timing().last_call_ms)
            line[f"n{n}"] = {"ms": min(ts), "Gint_s": n * n / (min(ts) * 1e-3) / 1e9}
        res[f"fp{prec}/v{var}"] = line
        print(f"fp{prec} v{var} err {err:.2e} " + " ".join(
            f"n={n}: {line[f'n{n}']['ms']:.3f} ms {line[f'n{n}']['Gint_s']:.1f} G/s" for n in ns), flush=True)
        var += 1
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "sweep.json"), "w"), indent=1)

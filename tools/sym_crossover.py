"""Where does the pair-symmetric kernel overtake the one-sided kernel, and with which build?
One RHS per size, kernel, variant and precision (CUDA-event time of nbody_rhs: force launch + K2),
best of 4.
   python tools/sym_crossover.py [--sizes 8192,16384,...] [--variants64 -1,8,5,3] [--variants32 -1,1,3,2]"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from parallel_comp_labs_b200 import capi, generators

ap = argparse.ArgumentParser()
ap.add_argument("--sizes", default="8192,16384,20000,24576,32768,49152,65536,98304,131072,262144")
ap.add_argument("--variants64", default="-1")
ap.add_argument("--variants32", default="-1")
ap.add_argument("--precisions", default="64,32")
a = ap.parse_args()


def best_rhs(ctx):
    best = 1e30
    for _ in range(4):
        ctx.rhs()
        best = min(best, ctx.timing().last_call_ms)
    return best


for n in [int(x) for x in a.sizes.split(",")]:
    for dist in ("plummer", "uniform"):
        bodies, eps = ((generators.plummer_bodies(n, 3), 1e-3) if dist == "plummer"
                       else (generators.uniform_bodies(n, 3), 1e-6))
        for prec in [int(x) for x in a.precisions.split(",")]:
            row = {"n": n, "masses": "equal" if dist == "plummer" else "general", "precision": prec}
            with capi.Context(n, prec) as ctx:
                ctx.set_params(6.67e-11, eps)
                ctx.set_option("pair_symmetry", 0)
                ctx.upload_aos(bodies)
                row["K1"] = n * float(n) / best_rhs(ctx) / 1e6
                ctx.set_option("pair_symmetry", 1)
                for v in [int(x) for x in (a.variants64 if prec == 64 else a.variants32).split(",")]:
                    ctx.set_option("sym_variant", v)
                    ms = best_rhs(ctx)
                    key = "K1s_auto" if v < 0 else f"K1s_v{v}"
                    row[key] = n * float(n) / ms / 1e6
                    if v < 0:
                        row["auto_variant"] = int(ctx.timing().sym_variant)
            print(json.dumps(row), flush=True)

"""Where does the pair-symmetric kernel overtake the one-sided kernel?  One RHS per size, kernel
and precision (CUDA-event time of nbody_rhs: force launch + K2), best of 3.
   python tools/sym_crossover.py [--sizes 8192,16384,...] > gpurun_out/crossover.json"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from parallel_comp_labs_b200 import capi, generators

ap = argparse.ArgumentParser()
ap.add_argument("--sizes", default="8192,16384,20000,24576,32768,49152,65536,98304,131072,262144")
a = ap.parse_args()
out = []
for n in [int(x) for x in a.sizes.split(",")]:
    for dist in ("plummer", "uniform"):
        bodies, eps = ((generators.plummer_bodies(n, 3), 1e-3) if dist == "plummer"
                       else (generators.uniform_bodies(n, 3), 1e-6))
        for prec in (64, 32):
            row = {"n": n, "masses": "equal" if dist == "plummer" else "general", "precision": prec}
            for sym in (0, 1):
                with capi.Context(n, prec) as ctx:
                    ctx.set_params(6.67e-11, eps)
                    ctx.set_option("pair_symmetry", sym)
                    ctx.upload_aos(bodies)
                    best = 1e30
                    for _ in range(4):
                        ctx.rhs()
                        best = min(best, ctx.timing().last_call_ms)
                    row["K1s" if sym else "K1"] = n * float(n) / best / 1e6
                    if sym:
                        row["sym_variant"] = int(ctx.timing().sym_variant)
            row["K1s_over_K1"] = row["K1s"] / row["K1"]
            out.append(row)
            print(json.dumps(row), flush=True)

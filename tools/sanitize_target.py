"""Target program for compute-sanitizer (tools/sanitize.sh): drives every kernel of the library
once or twice at small N through the C ABI -- K1 (all variants, FP64/FP32, equal and general
masses), K1s (all variants), K2 (RK2/RK4), K2g (general tableau, embedded step + undo), energy /
moments, snapshot pack, AoS/SoA pack + unpack.  No torch import (fast start under the tool)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from parallel_comp_labs_b200 import capi, generators, tableaux

sizes = [int(s) for s in (sys.argv[1:] or ["257", "3000"])]
for n in sizes:
    uni = generators.uniform_bodies(n, seed=11)          # general masses
    plu = generators.plummer_bodies(n, seed=12)          # equal masses
    if n > 1:
        plu[1, 1:4] = plu[0, 1:4] + 1e-5                   # inside the clamp radius: redo path
    for prec in (64, 32):
        for bodies, eps in ((uni, 1e-6), (plu, 1e-3)):
            with capi.Context(n, prec) as ctx:
                ctx.set_params(6.67e-11, eps)
                ctx.upload_aos(bodies)
                v = 0
                while True:                               # K1, every variant
                    try:
                        ctx.set_option("variant", v)
                    except capi.NBodyError:
                        break
                    ctx.rhs()
                    v += 1
                ctx.set_option("variant", -1)
                ctx.set_option("pair_symmetry", 1)
                s = 0
                while True:                               # K1s, every variant
                    try:
                        ctx.set_option("sym_variant", s)
                    except capi.NBodyError:
                        break
                    ctx.upload_aos(bodies)                # the fixed-point builds take their window from the upload
                    ctx.rhs()
                    s += 1
                ctx.set_option("sym_variant", -1)
                ctx.step(capi.NBODY_RK4, 1e-3, 1, sync=True)   # K1s + K2
                ctx.set_option("pair_symmetry", 0)
                ctx.set_option("use_graph", 0)
                ctx.step(capi.NBODY_RK4, 1e-3, 1, sync=True)   # K1 + K2
                ctx.step(capi.NBODY_RK2, 1e-3, 1, sync=True)
                ctx.set_option("use_graph", 1)
                ctx.step(capi.NBODY_RK4, 1e-3, 2, sync=True)   # graph replay
                tb = tableaux.TABLEAUX["dopri5"]
                ctx.step_rk(tb["a"], tb["b"], 1e-3, 1, sync=True)      # K2g
                ctx.step_embedded(tb["a"], tb["b"], tb["b_err"], 1e-3)
                ctx.step_undo()
                ctx.energy()
                ctx.snapshot()
                ctx.download_aos()
                print(f"n={n} fp{prec} {'general' if bodies is uni else 'equal'} masses: "
                      f"{v} K1 variants, {s} K1s variants ok", flush=True)
print("SANITIZE TARGET DONE")

"""Run a few RHS evaluations (for ncu captures).  usage: rhs_once.py prec n variant reps"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from parallel_comp_labs_b200 import capi, generators
prec, n, var, reps = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
bodies = generators.plummer_bodies(n, 3)
with capi.Context(n, prec) as ctx:
    ctx.set_params(6.67e-11, 1e-3)
    ctx.set_option("variant", var)
    ctx.upload_aos(bodies)
    for r in range(reps):
        ctx.rhs()
        print("rhs ms", ctx.timing().last_call_ms, "G/s", n * n / ctx.timing().last_call_ms / 1e6)

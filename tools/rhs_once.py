"""Run a few RHS evaluations (for ncu captures).
usage: rhs_once.py [--precision 64] [--bodies N] [--variant V] [--reps R]
                   [--pair-symmetry 1] [--sym-variant V] [--unimass -1|0]
   or the old positional form: rhs_once.py prec n variant reps"""
import argparse, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from parallel_comp_labs_b200 import capi, generators

ap = argparse.ArgumentParser()
ap.add_argument("positional", nargs="*", type=int)
ap.add_argument("--precision", type=int, default=64)
ap.add_argument("--bodies", type=int, default=131072)
ap.add_argument("--variant", type=int, default=-1)
ap.add_argument("--reps", type=int, default=2)
ap.add_argument("--pair-symmetry", type=int, default=0)
ap.add_argument("--sym-variant", type=int, default=-1)
ap.add_argument("--unimass", type=int, default=-1)
a = ap.parse_args()
if len(a.positional) == 4:
    a.precision, a.bodies, a.variant, a.reps = a.positional
elif a.positional:
    ap.error("positional form takes exactly: prec n variant reps")

bodies = generators.plummer_bodies(a.bodies, 3)
with capi.Context(a.bodies, a.precision) as ctx:
    ctx.set_params(6.67e-11, 1e-3)
    ctx.set_option("variant", a.variant)
    if a.unimass >= 0:
        ctx.set_option("unimass", a.unimass)
    if a.pair_symmetry:
        ctx.set_option("pair_symmetry", 1)
        ctx.set_option("sym_variant", a.sym_variant)
    ctx.upload_aos(bodies)
    for r in range(a.reps):
        ctx.rhs()
        ms = ctx.timing().last_call_ms
        print("rhs ms", ms, "G/s", a.bodies * a.bodies / ms / 1e6)

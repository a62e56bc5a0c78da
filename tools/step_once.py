"""Target for `ncu -k regex:stage_kernel`: upload a Plummer sphere and take RK4 steps through the
C ABI, so that the fused stage kernel (K2) runs on the benchmark's state.
   python tools/step_once.py [--bodies 1048576] [--precision 64] [--steps 1]"""
import argparse, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from parallel_comp_labs_b200 import capi, generators

ap = argparse.ArgumentParser()
ap.add_argument("--bodies", type=int, default=1 << 20)
ap.add_argument("--precision", type=int, default=64)
ap.add_argument("--steps", type=int, default=1)
a = ap.parse_args()
bodies = generators.plummer_bodies(a.bodies, 3)
with capi.Context(a.bodies, a.precision) as ctx:
    ctx.set_params(6.67e-11, 1e-3)
    ctx.upload_aos(bodies)
    ctx.step(capi.NBODY_RK4, 1e-3, a.steps, sync=True)
    t = ctx.timing()
    print("step ms", t.last_call_ms, "launches", t.total_launches, "sym_variant", t.sym_variant)

"""RHS time at N = 1M for forced j-slice counts (tail-length experiment)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from parallel_comp_labs_b200 import capi, generators
n = 1 << 20
bodies = generators.plummer_bodies(n, 0x5EED0004)
for s in [int(a) for a in sys.argv[1:]] or [0]:
    with capi.Context(n, 64) as ctx:
        ctx.set_params(6.67e-11, 1e-3)
        ctx.set_option("slices", s)
        ctx.upload_aos(bodies)
        ts = []
        for _ in range(3):
            ctx.rhs(); ts.append(ctx.timing().last_call_ms)
    print(f"slices={s}: {min(ts):.2f} ms  {n*n/min(ts)/1e6:.1f} G/s", flush=True)

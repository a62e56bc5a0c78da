"""Per-step latency of the reference's 4-body test (RK2, 200 steps), FP32 and FP64, with and
without CUDA-graph replay.  The reference's own published number for this run is
1.64925e-4 s/step (lab_5_N_body_CUDA/output/res_4body_float_64_.csv:2, unknown GPU)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from parallel_comp_labs_b200 import capi
b4 = np.array([[8810324116.227, 1, 0, 0, 0, .9, 0], [8810324116.227, 0, 1, 0, -.9, 0, 0],
               [8810324116.227, -1, 0, 0, 0, -.9, 0], [8810324116.227, 0, -1, 0, .9, 0, 0]])
for prec in (32, 64):
    for graph in (0, 1):
        with capi.Context(4, prec) as ctx:
            ctx.set_option("use_graph", graph)
            ctx.upload_aos(b4)
            ctx.step(capi.NBODY_RK2, 0.1, 20, sync=True)
            t0 = time.perf_counter()
            for _ in range(200):
                ctx.step(capi.NBODY_RK2, 0.1, 1)      # one call per step, as the host loop does
            ctx.sync()
            wall = (time.perf_counter() - t0) / 200
            ctx.step(capi.NBODY_RK2, 0.1, 200, sync=True)
            dev = ctx.timing().last_call_ms / 200 * 1e-3
        print(f"fp{prec} graph={graph}: {wall*1e6:.1f} us/step (host loop, wall)  {dev*1e6:.1f} us/step (device, 200 steps in one call)")

"""nbody_rhs_noalloc against the context-based nbody_rhs at the largest size the former takes:
device time per call (torch CUDA events on the default stream vs nbody_timing.last_call_ms).
   python tools/noalloc_timing.py"""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from parallel_comp_labs_b200 import capi, generators

out = []
for prec in (64, 32):
    n = capi.NOALLOC_MAX_BODIES
    bodies = generators.plummer_bodies(n, 3)
    dt = torch.float64 if prec == 64 else torch.float32
    m = torch.tensor(bodies[:, 0], dtype=dt, device="cuda:0")
    r4 = torch.zeros((n, 4), dtype=dt, device="cuda:0")
    r4[:, :3] = torch.tensor(bodies[:, 1:4], dtype=dt, device="cuda:0")
    acc = torch.empty((n, 3), dtype=dt, device="cuda:0")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ms = []
    for k in range(8):
        e0.record()
        capi.rhs_noalloc(n, m.data_ptr(), r4.data_ptr(), acc.data_ptr(), precision=prec, eps=1e-3)
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    with capi.Context(n, prec) as ctx:
        ctx.set_params(6.67e-11, 1e-3)
        ctx.set_option("unimass", 0)            # general-mass kernels, as the no-alloc entry runs
        ctx.set_option("pair_symmetry", 1)
        ctx.upload_aos(bodies)
        cms = []
        for k in range(8):
            want = ctx.rhs()
            cms.append(ctx.timing().last_call_ms)
        sv = ctx.timing().sym_variant
    got = acc.cpu().numpy().astype(np.float64)
    err = float(np.abs(got - want).max() / np.abs(want).max())
    out.append({"precision": prec, "n": n, "noalloc_ms": float(np.median(ms[3:])),
                "noalloc_G_int_per_s": n * n / float(np.median(ms[3:])) / 1e6,
                "context_ms": float(np.median(cms[3:])), "context_sym_variant": int(sv),
                "context_G_int_per_s": n * n / float(np.median(cms[3:])) / 1e6, "rel_err": err})
print(json.dumps(out))

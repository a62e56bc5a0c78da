import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from parallel_comp_labs_b200 import capi, generators
g = dict(np.load(os.path.join(ROOT, "tests/golden/golden.npz")))
def run(b, scheme, tau, steps, eps, graph, um):
    with capi.Context(b.shape[0], 64) as ctx:
        ctx.set_params(6.67e-11, eps); ctx.set_option("use_graph", graph); ctx.set_option("unimass", um)
        ctx.upload_aos(b); ctx.step(scheme, tau, steps, sync=True); return ctx.download_aos()
for name, b, want, tau, steps, eps in (("4body", g["body4"], g["ref_rk4_4body_0100_final"], 0.1, 200, 1e-6),
                                       ("plummer1000", g["plummer1000_in"], g["ref_rk4_plummer1000_3steps"], 1e-3, 3, 1e-3)):
    for graph in (0, 1):
        for um in (0, -1):
            got = run(b, 4, tau, steps, eps, graph, um)
            print(name, "graph", graph, "um", um, "err r", np.abs(got[:,1:4]-want[:,1:4]).max()/np.abs(want[:,1:4]).max(),
                  "err v", np.abs(got[:,4:7]-want[:,4:7]).max()/np.abs(want[:,4:7]).max(), flush=True)

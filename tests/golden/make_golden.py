"""Regenerate tests/golden/*.npz from the REFERENCE ITSELF (run in the build container,
where /root/reference exists and oracle/_ref has been built from it).

Everything stored here is either (a) output of the reference's own lab_4 code
(oracle/_ref/libref_lab4.so = the reference sources compiled where they lie) on committed
inputs, or (b) the reference's own known-answer files re-packed as arrays:
  lab_4_N_body_MPI/src/include/N_body/N_body_config/4body.txt, traj{1..4}.csv
  lab_4_N_body_MPI/output/rand_bodies.txt (first 512 bodies only)
  lab_5_N_body_CUDA/output/res4body_async.csv
The GPU box has no /root/reference; tests read only these fixtures there.
usage: python tests/golden/make_golden.py
"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import RefLab4, read_csv_rows  # noqa: E402
from parallel_comp_labs_b200 import generators  # noqa: E402

REF = "/root/reference"
L4 = REF + "/lab_4_N_body_MPI"
ref = RefLab4()
out = {}

# ---- the reference's own known-answer data --------------------------------------------------
b4 = ref.read_data(L4 + "/src/include/N_body/N_body_config/4body.txt")
out["body4"] = b4
for k in range(1, 5):
    rows = np.array([[float(x) for x in line.strip().split(";")]
                     for line in open(L4 + f"/src/include/N_body/N_body_config/traj{k}.csv").readlines()[1:]])
    out[f"traj{k}"] = rows  # (201, 4): time, r1, r2, r3
_, rows = read_csv_rows(REF + "/lab_5_N_body_CUDA/output/res4body_async.csv")
out["lab5_res4body_async"] = rows  # (200, 13) FP32 GPU trajectory of the reference
rb = ref.read_data(L4 + "/output/rand_bodies.txt")
out["rand512"] = rb[:512].copy()

# ---- reference RK2 on the 4-body problem (every row, lossless %.16e) ----------------------------
with tempfile.TemporaryDirectory() as td:
    for tau, key in ((0.1, "0100"), (0.05, "0050"), (0.025, "0025")):
        p = os.path.join(td, f"r{key}.csv")
        ref.rk2_csv(b4, tau, 0.0, 20.0, p, use_bodyr=True)
        header, rows = read_csv_rows(p)
        out[f"ref_rk2_4body_{key}"] = rows if key == "0100" else rows[::8]
        out[f"ref_rk2_4body_{key}_nrows"] = np.array([rows.shape[0]])
    out["ref_header"] = np.frombuffer(header.encode(), dtype=np.uint8)
    # byte-level sample of write_bodies_coords
    p = os.path.join(td, "w.csv")
    ref.write_bodies_coords(p, 0.30000000000000004, b4)
    out["ref_row_bytes"] = np.frombuffer(open(p, "rb").read(), dtype=np.uint8)
    # reference RK2 on the reference's random bodies (first 512), 11 steps, 3 thread-ranks
    p = os.path.join(td, "rand.csv")
    ref.rk2_csv(rb[:512], 0.1, 0.0, 1.0, p, n_ranks=3)
    _, rows = read_csv_rows(p)
    out["ref_rk2_rand512_rows"] = rows[[1, 6, 11]]  # after steps 1, 6, 11 (row 0 is t_start)
    out["ref_rk2_rand512_nrows"] = np.array([rows.shape[0]])

# ---- reference RHS ----------------------------------------------------------------------------
out["ref_rhs_body4"] = ref.rhs(b4)[:, 4:7]
out["ref_rhs_rand512"] = ref.rhs(rb[:512])[:, 4:7]
uni = generators.uniform_bodies(777, seed=0x5EED0001)  # ragged size on purpose
out["ref_rhs_uniform777"] = ref.rhs(uni)[:, 4:7]
plu = generators.plummer_bodies(1000, seed=0x5EED0003)
plu[10, 1:4] = plu[11, 1:4] + 2e-4  # a pair inside the clamp radius eps=1e-3
plu[20, 1:4] = plu[21, 1:4]          # coincident bodies
out["plummer1000_in"] = plu
out["ref_rhs_plummer1000_eps1e-3"] = ref.rhs(plu, eps=1e-3)[:, 4:7]

# ---- RK4 over the reference's RHS (NOT a reference integrator; see oracle/ref_lab4.cpp) --------
f4, n4 = ref.rk4(b4, 0.1, 0.0, 20.0)
out["ref_rk4_4body_0100_final"] = f4
out["ref_rk4_4body_0100_steps"] = np.array([n4])
fu, _ = ref.rk4(uni, 0.1, 0.0, 1.0, max_steps=3)
out["ref_rk4_uniform777_3steps"] = fu
fp, _ = ref.rk4(plu, 1e-3, 0.0, 1.0, eps=1e-3, max_steps=3)
out["ref_rk4_plummer1000_3steps"] = fp

np.savez_compressed(os.path.join(HERE, "golden.npz"), **out)
# the reference's 10 000-body input: the text holds 6 significant digits, so it is stored
# exactly as decimal mantissa (int32) x 10^(exponent-5); plus reference RHS rows of it.
man = np.zeros(rb.shape, dtype=np.int32)
ex = np.zeros(rb.shape, dtype=np.int8)
for idx, x in np.ndenumerate(rb):
    m, e = ("%.5e" % x).split("e")
    man[idx] = int(round(float(m) * 1e5))
    ex[idx] = int(e)
back = np.array([float("%de%d" % (m, e - 5)) for m, e in zip(man.ravel(), ex.ravel())]).reshape(rb.shape)
assert np.array_equal(back, rb)
np.savez_compressed(os.path.join(HERE, "rand10k.npz"), mantissa=man, exponent=ex,
                    ref_rhs_rows32=ref.rhs(rb, 0, 32)[:32, 4:7],
                    ref_rhs_rows_last8=ref.rhs(rb, 9992, 10000)[9992:, 4:7])
print({k: v.shape for k, v in out.items()})
print("golden.npz bytes:", os.path.getsize(os.path.join(HERE, "golden.npz")))

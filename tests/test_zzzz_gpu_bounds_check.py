"""Device-side bounds asserts instead of compute-sanitizer (closed on this GPU pool, see
profiles/r02_sanitizer.md): libnbody_b200_check.so is the product source compiled with
-DNBODY_BOUNDS_CHECK -- every global store / RED / bulk-copy source of K1 and K1s checks its index
against the allocation and traps.  tools/sanitize_target.py drives every kernel variant at ragged
small sizes through it; a trap surfaces as a CUDA launch failure -> NBodyError -> exit code != 0."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(900, method="thread")]

CHECK_LIB = os.path.join(ROOT, "parallel_comp_labs_b200", "libnbody_b200_check.so")


def _run(args):
    if not os.path.exists(CHECK_LIB):
        subprocess.run(["make", "-C", os.path.join(ROOT, "parallel_comp_labs_b200", "csrc"), "check"],
                       check=True, stdout=subprocess.DEVNULL)
    return subprocess.run([sys.executable] + args, capture_output=True, text=True, timeout=800,
                          env=dict(os.environ, NBODY_B200_LIB=CHECK_LIB))


def test_every_kernel_variant_under_bounds_asserts():
    r = _run([os.path.join(ROOT, "tools", "sanitize_target.py"), "1", "257", "3000", "5121"])
    assert r.returncode == 0 and "SANITIZE TARGET DONE" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.parametrize("prec", [64, 32])
def test_auto_selected_kernels_at_a_ragged_mid_size(prec):
    # 70 001 bodies: above the automatic pair-symmetry threshold, last block and last tile ragged
    r = _run([os.path.join(ROOT, "tools", "rhs_once.py"), "--precision", str(prec), "--bodies", "70001"])
    assert r.returncode == 0 and "rhs ms" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]

import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")
    # build the product library / driver and the checkers if this checkout has not been built
    need = [os.path.join(ROOT, "parallel_comp_labs_b200", "libnbody_b200.so"),
            os.path.join(ROOT, "parallel_comp_labs_b200", "bin", "nbody_b200"),
            os.path.join(ROOT, "oracle", "_build", "liboracle.so")]
    if not all(os.path.exists(f) for f in need):
        import subprocess
        subprocess.run(["make", "-C", os.path.join(ROOT, "parallel_comp_labs_b200", "csrc")],
                       check=True, stdout=subprocess.DEVNULL)
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle")], check=True,
                       stdout=subprocess.DEVNULL)


def _has_gpu() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return os.path.exists("/dev/nvidia0")


HAS_GPU = _has_gpu()


def pytest_collection_modifyitems(config, items):
    # GPU tests never silently pass on a CPU box: they are skipped there, and on a GPU box
    # every one of them goes through libnbody_b200.so (no fallback exists to pass through).
    if HAS_GPU:
        return
    skip = pytest.mark.skip(reason="no GPU in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    return dict(np.load(os.path.join(ROOT, "tests", "golden", "golden.npz")))


@pytest.fixture(scope="session")
def rand10k():
    z = np.load(os.path.join(ROOT, "tests", "golden", "rand10k.npz"))
    man, ex = z["mantissa"], z["exponent"]
    bodies = np.array([float("%de%d" % (m, e - 5)) for m, e in zip(man.ravel(), ex.ravel())])
    return bodies.reshape(man.shape), z["ref_rhs_rows32"], z["ref_rhs_rows_last8"]


@pytest.fixture(scope="session")
def oracle():
    from oracle import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def reflab4():
    from oracle import RefLab4
    if not RefLab4.available():
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    return RefLab4()


def rel_inf(a, b):
    """Inf-norm of the difference relative to the inf-norm of the reference value (the
    repo's norm idiom, lab_4 .../utils/norm_difference.hpp:16-23, applied to the whole state)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))

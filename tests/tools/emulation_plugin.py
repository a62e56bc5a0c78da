"""pytest plugin (TEST INFRASTRUCTURE): run the `-m gpu` parity tests on a machine WITHOUT a GPU
against the emulated library (tests/tools/nbody_capi_emu.cpp = the unmodified nbody_capi.cu over
cuda_rt_emu.h / cuda_emu.h / fake_nccl.cpp).  A development aid for checking new host logic and
kernel source before GPU time is spent; it proves nothing about the hardware and is never used by
the driver's `pytest -m gpu` run, which goes through libnbody_b200.so on a real B200.

    PYTHONPATH=tests/tools python -m pytest -p emulation_plugin tests/test_gpu_parity.py -m gpu \
        -p no:timeout --deselect tests/test_gpu_parity.py::test_rhs_large_n_row_subsample \
        --deselect tests/test_gpu_parity.py::test_rk2_rk4_uniform20k_config

(the two deselected cases are sized for a GPU: 262 144 and 20 000 bodies; in
tests/test_zzz_gpu_pair_symmetry.py deselect test_sym_large_n_row_subsample_and_momentum for the
same reason).  Tests that start the
compiled driver binary (tests/test_gpu_host_driver.py) or torchrun (tests/test_gpu_multi.py) need
real devices and are not covered; tests/test_capi_emulation.py holds their emulated counterparts.
EMU_SCHED_SEED=<n> randomises the order in which the emulated threads are visited."""
import ctypes as C
import os
import subprocess
import sys
import tempfile

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    d = tempfile.mkdtemp(prefix="nbody_emu")
    lib, fake = os.path.join(d, "libnbody_emu.so"), os.path.join(d, "libfake_nccl.so")
    tools = os.path.join(ROOT, "tests", "tools")
    subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-Wno-unknown-pragmas", "-w",
                    "-shared", "-fPIC", "-o", lib, os.path.join(tools, "nbody_capi_emu.cpp"),
                    "-ldl", "-lpthread"], check=True)
    subprocess.run(["g++", "-O1", "-std=c++17", "-shared", "-fPIC", "-o", fake,
                    os.path.join(tools, "fake_nccl.cpp"), "-lpthread"], check=True)
    os.environ["NBODY_NCCL_LIB"] = fake
    from parallel_comp_labs_b200 import capi
    capi._lib = capi._bind(C.CDLL(lib))
    print(f"[emulation_plugin] `-m gpu` tests run against {lib} (CPU emulation, not a GPU result)")


@pytest.hookimpl(tryfirst=True)
def pytest_collection_modifyitems(config, items):
    import conftest
    conftest.HAS_GPU = True   # do not skip the gpu-marked tests: they talk to the emulated library

"""parallel_comp_labs_b200 -- B200-native (sm_100a) rebuild of ONE hot path of
sanonesan/parallel_comp_labs: the all-pairs clamped-softening acceleration inside the
explicit Runge-Kutta N-body integrator (lab_4_N_body_MPI / lab_5_N_body_CUDA).

Layout: ``csrc/`` CUDA kernels + the C ABI (``include/nbody_b200.h``), ``host/`` the C++
host driver mirroring the lab's interface, ``capi.py``/``lab.py`` the ctypes binding and
the Python mirror of the same interface used by tests and bench, ``generators.py`` the
synthetic inputs.  No CPU implementation of the path lives in this package.
"""
from . import capi, generators, lab  # noqa: F401
from .capi import NBODY_RK2, NBODY_RK4, Context, NBodyError  # noqa: F401

__all__ = ["capi", "generators", "lab", "Context", "NBodyError", "NBODY_RK2", "NBODY_RK4"]

"""The product's kernel SOURCE executed on the CPU.

tests/tools/kernel_emu.cpp compiles parallel_comp_labs_b200/csrc/{force,force_sym,stage}_kernels.cuh
with g++ against tests/tools/cuda_emu.h (CUDA threads as fibers; real rendezvous for
__syncthreads / __shfl_sync / __any_sync; the mbarrier + bulk-copy ring with deadlock and
write-after-read checks; MUFU seeds modelled at their measured accuracy) and runs
pack -> K1 or K1s -> K2 exactly as nbody_rhs launches them on one GPU.  This is a check of the
kernels' logic (index arithmetic, shared-memory layout, ring protocol, clamp redo, rotation,
slice reduction) that needs no GPU; the GPU parity suite (-m gpu) remains the proof of the bits.
Nothing here is a product path: the library has no CPU fallback."""
import ctypes as C
import os
import subprocess
import tempfile

import numpy as np
import pytest

from conftest import ROOT, rel_inf
from parallel_comp_labs_b200 import generators


@pytest.fixture(scope="module")
def emu_lib():
    out = os.path.join(tempfile.mkdtemp(prefix="kemu"), "libkemu.so")
    subprocess.run(["g++", "-O1", "-std=c++17", "-ffp-contract=off", "-Wno-unknown-pragmas",
                    "-shared", "-fPIC", "-o", out,
                    os.path.join(ROOT, "tests", "tools", "kernel_emu.cpp")], check=True)
    L = C.CDLL(out)
    L.emu_rhs.argtypes = [C.c_uint, C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_int, C.c_int,
                          C.c_uint, C.c_int, C.c_uint, C.c_void_p, C.c_char_p, C.c_int]
    L.emu_rhs_sym_ranks.argtypes = [C.c_uint, C.c_void_p, C.c_double, C.c_double, C.c_uint, C.c_uint,
                                    C.c_int, C.c_void_p, C.c_char_p, C.c_int]
    return L


@pytest.fixture(scope="module")
def emu(emu_lib):
    L = emu_lib

    def run(bodies, prec=64, eps=1e-6, G=6.67e-11, sym=0, ipt=2, slices=1, unimass=1, grid=2):
        b = np.ascontiguousarray(bodies, dtype=np.float64)
        acc = np.zeros((b.shape[0], 3))
        err = C.create_string_buffer(256)
        rc = L.emu_rhs(b.shape[0], b.ctypes.data, prec, G, eps, sym, ipt, slices, unimass, grid,
                       acc.ctypes.data, err, 256)
        assert rc == 0, (rc, err.value.decode())
        return acc
    return run


# (sym, i-bodies per thread): the default kernel's two register blockings and every blocking the
# pair-symmetric kernel is instantiated with
MODES = [(0, 2), (0, 4), (1, 2), (1, 4), (1, 6), (1, 8)]


@pytest.mark.parametrize("sym,ipt", MODES)
def test_emulated_kernels_ragged_sizes(emu, oracle, golden, sym, ipt):
    for n, slices, grid in ((1, 1, 1), (4, 1, 1), (257, 2, 3), (700, 3, 2), (1100, 2, 4)):
        b = golden["body4"] if n == 4 else generators.uniform_bodies(n, seed=50 + n)
        want = oracle.rhs(b)[:, 4:7]
        got = emu(b, sym=sym, ipt=ipt, slices=slices, grid=grid)
        if n == 1:
            assert np.all(got == 0)
        else:
            assert rel_inf(got, want) < 1e-14, (n, slices, grid)


@pytest.mark.parametrize("sym,ipt", [(0, 2), (1, 4), (1, 6)])
def test_emulated_kernels_clamp_masses_and_phantom_tiles(emu, oracle, sym, ipt):
    """Plummer sphere with a pair inside the clamp radius, a coincident pair in different blocks
    (the unclamped pass produces NaN there: only the redo makes it finite), general masses with a
    massless body; 11 j-tiles, so the pair-symmetric walk names a phantom tile."""
    n = 2600
    b = generators.plummer_bodies(n, seed=5)
    b[7, 1:4] = b[8, 1:4] + 1e-5
    b[1500, 1:4] = b[20, 1:4]
    b[2100, 1:4] = b[600, 1:4] + 2e-4
    want = oracle.rhs(b, eps=1e-3, threads=8)[:, 4:7]
    got = emu(b, eps=1e-3, sym=sym, ipt=ipt, slices=2, unimass=1, grid=3)
    assert np.all(np.isfinite(got)) and rel_inf(got, want) < 2e-14
    g = b.copy()
    g[:, 0] *= np.linspace(0.5, 2.0, n)
    g[33, 0] = 0.0
    want = oracle.rhs(g, eps=1e-3, threads=8)[:, 4:7]
    got = emu(g, eps=1e-3, sym=sym, ipt=ipt, slices=2, unimass=0, grid=3)
    assert np.all(np.isfinite(got)) and rel_inf(got, want) < 2e-14


@pytest.mark.parametrize("sym,ipt", [(0, 2), (1, 4)])
def test_emulated_kernels_eps_extremes_and_fp32(emu, oracle, sym, ipt):
    u = generators.uniform_bodies(600, seed=3)
    # eps = 0: NaN from the self pair, as in the reference (struct_ode_system.hpp:69-72)
    assert np.all(np.isnan(emu(u, eps=0.0, sym=sym, ipt=ipt, unimass=0)))
    want = oracle.rhs(u, eps=1e4)[:, 4:7]          # every pair inside the clamp radius
    assert rel_inf(emu(u, eps=1e4, sym=sym, ipt=ipt, unimass=0), want) < 1e-14
    want = oracle.rhs(u)[:, 4:7]
    got = emu(u, prec=32, sym=sym, ipt=ipt, unimass=0)
    assert rel_inf(got, want) < 1e-4               # north_star's FP32 tolerance


def test_emulated_pair_symmetric_kernel_on_several_ranks(emu_lib, oracle):
    """world > 1: every rank walks its own blocks of the global schedule (block0, nb_own) and
    collects reactions for all bodies; summing the j-side arrays over the ranks (the
    reduce-scatter) and reducing the owned rows in K2 gives the reference operator."""
    L = emu_lib
    for n, world, slices, um in ((1000, 2, 1, 0), (1500, 3, 2, 1), (1900, 2, 2, 1)):
        b = generators.uniform_bodies(n, seed=n)
        if um:
            b[:, 0] = 2e7
        acc = np.full((n, 3), np.nan)
        err = C.create_string_buffer(256)
        rc = L.emu_rhs_sym_ranks(n, b.ctypes.data, 6.67e-11, 1e-6, world, slices, um,
                                 acc.ctypes.data, err, 256)
        assert rc == 0, (rc, err.value.decode())
        assert rel_inf(acc, oracle.rhs(b)[:, 4:7]) < 1e-14


def test_emulation_catches_deliberate_defects():
    """The checker checked: a ring kernel built on the product's mbarrier / bulk-copy helpers runs
    clean, and its three defective twins are reported -- a stage refilled before every reader has
    passed a barrier, a lane that skips a warp collective, a wait on a stage that was never armed."""
    out = os.path.join(tempfile.mkdtemp(prefix="emuself"), "libemuself.so")
    subprocess.run(["g++", "-O1", "-std=c++17", "-Wno-unknown-pragmas", "-shared", "-fPIC", "-o", out,
                    os.path.join(ROOT, "tests", "tools", "emu_selftest.cpp")], check=True)
    L = C.CDLL(out)
    L.selftest.restype = C.c_char_p
    assert L.selftest(0) == b""
    assert b"may still be reading" in L.selftest(1)
    assert b"deadlock" in L.selftest(2)
    assert b"deadlock" in L.selftest(3)

"""The fixed-point row sums behind option "deterministic" (parallel_comp_labs_b200/csrc/fixed_sums.cuh)
against exact rational arithmetic: the SOURCE the GPU runs, compiled for the host
(tests/tools/fx_sums_host.cpp).  What is checked is what the mode promises: the limbs do not
depend on the order of the additions, and the value read back is the correctly rounded exact sum."""
import ctypes as C
import os
import subprocess
import tempfile
from fractions import Fraction

import numpy as np
import pytest

from conftest import ROOT


@pytest.fixture(scope="module")
def fx():
    out = os.path.join(tempfile.mkdtemp(prefix="fxsums"), "libfx.so")
    subprocess.run(["g++", "-O1", "-std=c++17", "-ffp-contract=off", "-Wno-unknown-pragmas", "-shared",
                    "-fPIC", "-o", out, os.path.join(ROOT, "tests", "tools", "fx_sums_host.cpp")],
                   check=True)
    L = C.CDLL(out)
    L.fx_accumulate.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p]
    L.fx_value.argtypes = [C.c_void_p, C.c_int]
    L.fx_value.restype = C.c_double
    return L


def accumulate(L, x, e0):
    x = np.ascontiguousarray(x, dtype=np.float64)
    limbs = np.zeros(L.fx_limbs(), dtype=np.uint64)
    poison = np.zeros(1, dtype=np.uint64)
    L.fx_accumulate(x.ctypes.data, x.size, e0, limbs.ctypes.data, poison.ctypes.data)
    return limbs, int(poison[0])


def exact_truncated(x, e0):
    """Sum of the values with the bits below 2^e0 dropped towards zero, as a Fraction."""
    unit = Fraction(2) ** e0
    tot = Fraction(0)
    for v in x:
        q = Fraction(float(v)) / unit
        tot += (abs(q).numerator // abs(q).denominator) * (1 if q >= 0 else -1)
    return tot * unit


def test_order_independent_and_correctly_rounded(fx):
    rng = np.random.default_rng(7)
    bits = 32 * fx.fx_limbs()
    for trial in range(40):
        n = int(rng.integers(1, 400))
        top = int(rng.integers(-300, 300))            # window [2^(top-bits), 2^top)
        e0 = top - bits
        # magnitudes spread over up to 120 binades under the top, mixed signs, heavy cancellation
        ex = top - 2 - rng.integers(0, 120, size=n)
        x = np.ldexp(rng.uniform(0.5, 1.0, size=n), ex) * rng.choice([-1.0, 1.0], size=n)
        if trial % 3 == 0:                            # exact cancellation down to the small terms
            x = np.concatenate([x, -x[: n // 2]])
        limbs, poison = accumulate(fx, x, e0)
        assert poison == 0
        for _ in range(3):
            perm = rng.permutation(x.size)
            l2, p2 = accumulate(fx, x[perm], e0)
            assert p2 == 0 and np.array_equal(limbs, l2)          # bit-identical whatever the order
        want = exact_truncated(x, e0)
        got = fx.fx_value(limbs.ctypes.data, e0)
        assert got == float(want)                                 # one correct rounding of the exact sum
        # nothing was truncated here (every term has its 53 bits inside the window) ...
        assert want == sum(Fraction(float(v)) for v in x)


def test_bits_below_the_window_are_dropped_towards_zero(fx):
    e0 = -40
    x = np.array([1.0 + 2.0 ** -52, -(2.0 ** -41), 2.0 ** -41 + 2.0 ** -60, 3.0 * 2.0 ** -40, -(2.0 ** -100)])
    limbs, poison = accumulate(fx, x, e0)
    assert poison == 0
    assert fx.fx_value(limbs.ctypes.data, e0) == float(exact_truncated(x, e0))
    assert exact_truncated(x, e0) == Fraction(1) + 3 * Fraction(2) ** -40   # only these bits are in the window


def test_zero_negative_total_and_single_bit(fx):
    e0 = -100
    for x in ([0.0, -0.0], [1.5, -1.5], [-3.25], [2.0 ** -100], [-(2.0 ** -100)], [1e-20, -1e-3, 7.0, -7.0]):
        limbs, poison = accumulate(fx, np.array(x), e0)
        assert poison == 0
        assert fx.fx_value(limbs.ctypes.data, e0) == float(exact_truncated(x, e0))
    # many equal terms: the deferred carries (31 bits of headroom per limb) are resolved at read time
    x = np.full(100000, 0.1)
    limbs, _ = accumulate(fx, x, e0)
    assert fx.fx_value(limbs.ctypes.data, e0) == float(exact_truncated(x, e0))


def test_non_finite_and_out_of_window_poison_the_row(fx):
    bits = 32 * fx.fx_limbs()
    for bad in (np.inf, -np.inf, np.nan):
        _, poison = accumulate(fx, np.array([1.0, bad, 2.0]), -100)
        assert poison == 1
    _, poison = accumulate(fx, np.array([2.0 ** (bits - 100)]), -100)       # at the top of the window
    assert poison == 1
    limbs, poison = accumulate(fx, np.array([2.0 ** (bits - 101)]), -100)   # just inside
    assert poison == 0 and fx.fx_value(limbs.ctypes.data, -100) == 2.0 ** (bits - 101)


def test_property_random_lists_any_order_correct_rounding(fx):
    """Property test (hypothesis): arbitrary finite doubles spread over the whole window -- fast path,
    slow path and partly truncated values mixed -- sum to the same limbs in any order, and read back
    as the correctly rounded exact (truncated) sum."""
    from hypothesis import given, settings, strategies as st
    bits = 32 * fx.fx_limbs()

    @settings(max_examples=150, deadline=None)
    @given(st.integers(-400, 400), st.lists(st.tuples(st.floats(0.5, 1.0, exclude_max=True), st.integers(-230, -3),
                                                       st.booleans()), min_size=1, max_size=60), st.randoms())
    def check(top, terms, rnd):
        e0 = top - bits
        x = np.array([np.ldexp(m, top + ex) * (-1.0 if neg else 1.0) for m, ex, neg in terms])
        limbs, poison = accumulate(fx, x, e0)
        assert poison == 0
        perm = list(range(x.size))
        rnd.shuffle(perm)
        l2, _ = accumulate(fx, x[perm], e0)
        assert np.array_equal(limbs, l2)
        assert fx.fx_value(limbs.ctypes.data, e0) == float(exact_truncated(x, e0))

    check()

"""ctypes front-end of the CPU checkers (oracle/).  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package; the product
(``parallel_comp_labs_b200``) never does.

* :class:`Oracle`  -- ``oracle/_build/liboracle.so``, the plain-C restatement
  (``oracle/nbody_oracle.c``; every function cites the reference file:line).
* :class:`RefLab4` -- ``oracle/_ref/libref_lab4.so``, the reference's own lab_4
  sources compiled where they lie (``oracle/ref_lab4.cpp``), when that build exists.

Bodies are numpy arrays of shape (n, 7): ``m, x, y, z, vx, vy, vz`` -- the memory
image of ``Body<T>`` (lab_4_N_body_MPI/src/include/N_body/struct_Body.hpp:5-10).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "_build", "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libref_lab4.so")

_dp = C.POINTER(C.c_double)
_fp = C.POINTER(C.c_float)


def build(quiet: bool = True) -> None:
    """Compile the checkers (``make -C oracle``).  Building is not using."""
    subprocess.run(["make", "-C", HERE], check=True,
                   stdout=subprocess.DEVNULL if quiet else None)


def _aos(a, dtype):
    a = np.ascontiguousarray(a, dtype=dtype)
    assert a.ndim == 2 and a.shape[1] == 7, a.shape
    return a


class Oracle:
    def __init__(self):
        if not os.path.exists(ORACLE_SO):
            build()
        L = C.CDLL(ORACLE_SO)
        L.oracle_rhs_f64.argtypes = [_dp, _dp, C.c_size_t, C.c_size_t, C.c_size_t,
                                     C.c_double, C.c_double, C.c_int]
        L.oracle_rhs_f32.argtypes = [_fp, _fp, C.c_size_t, C.c_size_t, C.c_size_t,
                                     C.c_float, C.c_float, C.c_int]
        for nm in ("oracle_rk2_f64", "oracle_rk4_f64"):
            f = getattr(L, nm)
            f.restype = C.c_long
            f.argtypes = [_dp, C.c_size_t, C.c_double, C.c_double, C.c_double, C.c_double,
                          C.c_double, _dp, C.c_long, C.c_long, C.c_int]
        L.oracle_rk_subdiag_f64.restype = C.c_long
        L.oracle_rk_subdiag_f64.argtypes = [_dp, C.c_size_t, C.c_int, _dp, _dp, C.c_double,
                                            C.c_double, C.c_double, C.c_double, C.c_double,
                                            C.c_double, C.c_long, C.c_int]
        L.oracle_rk_general_f64.restype = C.c_long
        L.oracle_rk_general_f64.argtypes = [_dp, C.c_size_t, C.c_int, _dp, _dp, _dp, C.c_double,
                                            C.c_long, C.c_double, C.c_double, C.c_int, _dp]
        L.oracle_rk2_f32.restype = C.c_long
        L.oracle_rk2_f32.argtypes = [_fp, C.c_size_t, C.c_float, C.c_float, C.c_float,
                                     C.c_float, C.c_float, _fp, C.c_long, C.c_long, C.c_int]
        L.oracle_step_count_f64.restype = C.c_long
        L.oracle_step_count_f64.argtypes = [C.c_double] * 3
        L.oracle_step_count_f32.restype = C.c_long
        L.oracle_step_count_f32.argtypes = [C.c_float] * 3
        L.oracle_partition.argtypes = [C.c_size_t, C.c_int, C.POINTER(C.c_int),
                                       C.POINTER(C.c_int)]
        L.oracle_energy_f64.argtypes = [_dp, C.c_size_t, C.c_double, C.c_double, _dp, _dp,
                                        C.c_int]
        L.oracle_max_threads.restype = C.c_int
        self.L = L

    def max_threads(self) -> int:
        return int(self.L.oracle_max_threads())

    def rhs(self, bodies, begin=0, end=None, eps=1e-6, G=6.67e-11, threads=1):
        """diff rows [begin,end): returns (n,7) with .r = v and .v = acceleration."""
        b = _aos(bodies, np.float64)
        n = b.shape[0]
        end = n if end is None else end
        d = np.zeros_like(b)
        self.L.oracle_rhs_f64(b.ctypes.data_as(_dp), d.ctypes.data_as(_dp), n, begin, end,
                              eps, G, threads)
        return d

    def rhs_f32(self, bodies, begin=0, end=None, eps=1e-6, G=6.67e-11, threads=1):
        b = _aos(bodies, np.float32)
        n = b.shape[0]
        end = n if end is None else end
        d = np.zeros_like(b)
        self.L.oracle_rhs_f32(b.ctypes.data_as(_fp), d.ctypes.data_as(_fp), n, begin, end,
                              eps, G, threads)
        return d

    def _run(self, fn, bodies, tau, t_start, t_final, eps, G, max_steps, rows, threads,
             dtype, ptr):
        b = _aos(bodies, dtype).copy()
        n = b.shape[0]
        if dtype == np.float64:
            nsteps = self.step_count(t_start, t_final, tau)
        else:
            nsteps = self.step_count_f32(t_start, t_final, tau)
        if max_steps:
            nsteps = min(nsteps, max_steps)
        out = np.zeros((nsteps if rows else 0, 1 + 3 * n), dtype=dtype)
        done = fn(b.ctypes.data_as(ptr), n, tau, t_start, t_final, eps, G,
                  out.ctypes.data_as(ptr) if rows else None, nsteps if rows else 0,
                  max_steps or 0, threads)
        assert done == nsteps, (done, nsteps)
        return b, out

    def rk2(self, bodies, tau, t_start, t_final, eps=1e-6, G=6.67e-11, max_steps=0,
            rows=True, threads=1):
        """Reference RK2 loop; returns (final (n,7) state, rows [steps, 1+3n])."""
        return self._run(self.L.oracle_rk2_f64, bodies, tau, t_start, t_final, eps, G,
                         max_steps, rows, threads, np.float64, _dp)

    def rk4(self, bodies, tau, t_start, t_final, eps=1e-6, G=6.67e-11, max_steps=0,
            rows=True, threads=1):
        return self._run(self.L.oracle_rk4_f64, bodies, tau, t_start, t_final, eps, G,
                         max_steps, rows, threads, np.float64, _dp)

    def rk_subdiag(self, bodies, a, b, scale, tau, t_start, t_final, eps=1e-6, G=6.67e-11,
                   max_steps=0, threads=1):
        """Explicit RK with a sub-diagonal tableau (not in the reference); returns the final state."""
        x = _aos(bodies, np.float64).copy()
        a = np.ascontiguousarray(a, dtype=np.float64)
        b = np.ascontiguousarray(b, dtype=np.float64)
        assert b.size == a.size + 1
        self.L.oracle_rk_subdiag_f64(x.ctypes.data_as(_dp), x.shape[0], b.size,
                                     a.ctypes.data_as(_dp), b.ctypes.data_as(_dp), scale, tau,
                                     t_start, t_final, eps, G, max_steps, threads)
        return x

    def rk_general(self, bodies, a, b, tau, n_steps, b_err=None, eps=1e-6, G=6.67e-11, threads=1):
        """General explicit RK with the whole tableau as data (not in the reference): returns
        (final state, err[4] of the last step = max|err_r|, max|err_v|, max|r|, max|v|)."""
        x = _aos(bodies, np.float64).copy()
        a = np.ascontiguousarray(a, dtype=np.float64)
        b = np.ascontiguousarray(b, dtype=np.float64)
        s = b.size
        assert a.shape == (s, s)
        be = None if b_err is None else np.ascontiguousarray(b_err, dtype=np.float64)
        err = np.zeros(4)
        self.L.oracle_rk_general_f64(x.ctypes.data_as(_dp), x.shape[0], s, a.ctypes.data_as(_dp),
                                     b.ctypes.data_as(_dp),
                                     be.ctypes.data_as(_dp) if be is not None else None,
                                     tau, n_steps, eps, G, threads, err.ctypes.data_as(_dp))
        return x, err

    def rk2_f32(self, bodies, tau, t_start, t_final, eps=1e-6, G=6.67e-11, max_steps=0,
                rows=True, threads=1):
        return self._run(self.L.oracle_rk2_f32, bodies, tau, t_start, t_final, eps, G,
                         max_steps, rows, threads, np.float32, _fp)

    def step_count(self, t_start, t_final, tau) -> int:
        return int(self.L.oracle_step_count_f64(t_start, t_final, tau))

    def step_count_f32(self, t_start, t_final, tau) -> int:
        return int(self.L.oracle_step_count_f32(t_start, t_final, tau))

    def partition(self, n, n_procs):
        counts = (C.c_int * n_procs)()
        offsets = (C.c_int * n_procs)()
        rc = self.L.oracle_partition(n, n_procs, counts, offsets)
        if rc != 0:
            raise ValueError("The number of processes is greater than the problem dimension")
        return list(counts), list(offsets)

    def energy(self, bodies, eps=1e-6, G=6.67e-11, threads=1):
        b = _aos(bodies, np.float64)
        k = C.c_double()
        u = C.c_double()
        self.L.oracle_energy_f64(b.ctypes.data_as(_dp), b.shape[0], eps, G, C.byref(k),
                                 C.byref(u), threads)
        return k.value, u.value


class RefLab4:
    """The reference's own lab_4 code (oracle/_ref).  ``available()`` is False where
    neither /root/reference nor a prebuilt oracle/_ref exists."""

    @staticmethod
    def available() -> bool:
        return os.path.exists(REF_SO)

    def __init__(self):
        if not os.path.exists(REF_SO):
            raise FileNotFoundError(REF_SO)
        L = C.CDLL(REF_SO)
        L.ref_rhs.argtypes = [_dp, _dp, C.c_size_t, C.c_size_t, C.c_size_t, C.c_double,
                              C.c_double]
        L.ref_rhs_ranks.restype = C.c_double
        L.ref_rhs_ranks.argtypes = [_dp, _dp, C.c_size_t, C.c_size_t, C.c_int, C.c_double,
                                    C.c_double]
        L.ref_rk2_run.restype = C.c_double
        L.ref_rk2_run.argtypes = [_dp, C.c_size_t, C.c_double, C.c_double, C.c_double,
                                  C.c_double, C.c_double, C.c_int, C.c_int, C.c_char_p]
        L.ref_solve_parallel.restype = C.c_double
        L.ref_solve_parallel.argtypes = [_dp, C.c_size_t, C.c_double, C.c_double, C.c_double,
                                         C.c_double, C.c_double, C.c_int]
        L.ref_rk4.restype = C.c_long
        L.ref_rk4.argtypes = [_dp, C.c_size_t, C.c_double, C.c_double, C.c_double, C.c_double,
                              C.c_double, _dp, C.c_long, C.c_long]
        L.ref_read_data.restype = C.c_long
        L.ref_read_data.argtypes = [C.c_char_p, _dp, C.c_size_t]
        L.ref_write_bodies_coords.argtypes = [C.c_char_p, C.c_double, _dp, C.c_size_t]
        self.L = L

    def rhs(self, bodies, begin=0, end=None, eps=1e-6, G=6.67e-11):
        b = _aos(bodies, np.float64)
        n = b.shape[0]
        end = n if end is None else end
        d = np.zeros_like(b)
        rc = self.L.ref_rhs(b.ctypes.data_as(_dp), d.ctypes.data_as(_dp), n, begin, end, eps, G)
        assert rc == 0
        return d

    def rhs_ranks(self, bodies, rows_per_rank, n_ranks, eps=1e-6, G=6.67e-11, want_diff=False):
        """Rows [0, rows_per_rank*n_ranks) evaluated by n_ranks thread-ranks through the
        reference's N_body_func; returns (wall seconds, diff or None)."""
        b = _aos(bodies, np.float64)
        d = np.zeros_like(b) if want_diff else None
        w = self.L.ref_rhs_ranks(b.ctypes.data_as(_dp), d.ctypes.data_as(_dp) if want_diff else None,
                                 b.shape[0], rows_per_rank, n_ranks, eps, G)
        if w < 0:
            raise RuntimeError(f"ref_rhs_ranks failed ({w})")
        return w, d

    def rk2_csv(self, bodies, tau, t_start, t_final, csv_path, eps=1e-6, G=6.67e-11,
                n_ranks=1, use_bodyr=False) -> float:
        b = _aos(bodies, np.float64)
        w = self.L.ref_rk2_run(b.ctypes.data_as(_dp), b.shape[0], tau, t_start, t_final, eps, G,
                               n_ranks, int(use_bodyr), csv_path.encode())
        if w < 0:
            raise RuntimeError(f"ref_rk2_run failed ({w})")
        return w

    def solve_parallel(self, bodies, tau, t_start, t_final, eps=1e-6, G=6.67e-11,
                       n_ranks=1) -> float:
        b = _aos(bodies, np.float64)
        w = self.L.ref_solve_parallel(b.ctypes.data_as(_dp), b.shape[0], tau, t_start, t_final,
                                      eps, G, n_ranks)
        if w < 0:
            raise RuntimeError("ref_solve_parallel failed")
        return w

    def rk4(self, bodies, tau, t_start, t_final, eps=1e-6, G=6.67e-11, max_steps=0):
        b = _aos(bodies, np.float64).copy()
        n = b.shape[0]
        done = self.L.ref_rk4(b.ctypes.data_as(_dp), n, tau, t_start, t_final, eps, G, None, 0,
                              max_steps)
        assert done > 0
        return b, int(done)

    def read_data(self, path):
        n = self.L.ref_read_data(path.encode(), None, 0)
        if n < 0:
            raise IOError("file is't open")
        out = np.zeros((n, 7))
        self.L.ref_read_data(path.encode(), out.ctypes.data_as(_dp), n)
        return out

    def write_bodies_coords(self, path, time, bodies):
        b = _aos(bodies, np.float64)
        assert self.L.ref_write_bodies_coords(path.encode(), time, b.ctypes.data_as(_dp),
                                              b.shape[0]) == 0


def read_csv_rows(path):
    """Parse a reference-format trajectory CSV (';'-separated, trailing ';')."""
    rows = []
    with open(path) as f:
        header = f.readline()
        for line in f:
            parts = [p for p in line.strip().split(";") if p != ""]
            if parts:
                rows.append([float(p) for p in parts])
    return header, np.array(rows)

"""ctypes binding of the C ABI (include/nbody_b200.h -> libnbody_b200.so).

This is the only way Python reaches the device: there is no Python/CPU implementation
of the path behind it.  Importing works anywhere (the "does the library load and export
every symbol" check runs without a GPU); creating a context without an sm_100 device
raises :class:`NBodyError`.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG_DIR, "libnbody_b200.so")
HEADER_PATH = os.path.join(os.path.dirname(PKG_DIR), "include", "nbody_b200.h")

NBODY_RK2, NBODY_RK4 = 2, 4
NCCL_UID_BYTES = 128
P2P_HANDLE_BYTES = 256
SNAPSHOT_SLOTS = 4


class NBodyError(RuntimeError):
    pass


class Timing(C.Structure):
    _fields_ = [("last_call_ms", C.c_double), ("force_ms", C.c_double),
                ("force_launches", C.c_uint64), ("total_launches", C.c_uint64),
                ("interactions", C.c_uint64), ("steps_done", C.c_uint64),
                ("uniform_mass", C.c_uint32), ("variant", C.c_uint32),
                ("pair_evaluations", C.c_uint64), ("pair_symmetry", C.c_uint32),
                ("sym_variant", C.c_uint32), ("graph_captures", C.c_uint64),
                ("deterministic", C.c_uint32), ("reserved0", C.c_uint32)]


class TraceEntry(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("stream", C.c_uint32), ("start_ms", C.c_double),
                ("dur_ms", C.c_double)]


TRACE_KINDS = ("force", "stage", "memset", "all_gather", "reduce_scatter")


def build_library(quiet: bool = True) -> None:
    subprocess.run(["make", "-C", os.path.join(PKG_DIR, "csrc")], check=True,
                   stdout=subprocess.DEVNULL if quiet else None)


_lib = None


def lib() -> C.CDLL:
    """Load libnbody_b200.so (fails loudly if it has not been built)."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("NBODY_B200_LIB", LIB_PATH)   # e.g. the bounds-checked build (make check)
    if path != LIB_PATH:
        _lib = _bind(C.CDLL(path))
        return _lib
    if not os.path.exists(LIB_PATH):
        raise NBodyError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; "
                         f"g.build()'` (or make -C parallel_comp_labs_b200/csrc). "
                         f"There is no CPU fallback.")
    _lib = _bind(C.CDLL(LIB_PATH))
    return _lib


def _bind(L: C.CDLL) -> C.CDLL:
    """Declare the C signatures of include/nbody_b200.h on a loaded library."""
    vp, sz = C.c_void_p, C.c_size_t
    L.nbody_version.restype = C.c_char_p
    L.nbody_last_error.restype = C.c_char_p
    L.nbody_last_error.argtypes = [vp]
    L.nbody_nccl_unique_id.argtypes = [vp]
    L.nbody_device_count.restype = C.c_int
    L.nbody_create.argtypes = [C.POINTER(vp), sz, C.c_int, C.c_int, C.c_int, C.c_int, vp]
    L.nbody_destroy.argtypes = [vp]
    L.nbody_destroy.restype = None
    L.nbody_partition.argtypes = [sz, C.c_int, C.c_int, C.POINTER(sz), C.POINTER(sz),
                                  C.POINTER(sz)]
    L.nbody_p2p_export.argtypes = [vp, vp]
    L.nbody_p2p_connect.argtypes = [vp, vp]
    L.nbody_set_params.argtypes = [vp, C.c_double, C.c_double]
    L.nbody_set_option.argtypes = [vp, C.c_char_p, C.c_long]
    L.nbody_owned_range.argtypes = [vp, C.POINTER(sz), C.POINTER(sz)]
    L.nbody_upload_aos.argtypes = [vp, vp]
    L.nbody_upload_soa.argtypes = [vp, vp, vp, vp]
    L.nbody_upload_aos_owned.argtypes = [vp, vp]
    L.nbody_download_aos_owned.argtypes = [vp, vp]
    L.nbody_rhs.argtypes = [vp, vp]
    L.nbody_step.argtypes = [vp, C.c_int, C.c_double, C.c_long]
    L.nbody_step_tableau.argtypes = [vp, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double),
                                     C.c_double, C.c_double, C.c_long]
    L.nbody_step_rk.argtypes = [vp, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double),
                                C.c_double, C.c_long]
    L.nbody_step_embedded.argtypes = [vp, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double),
                                      C.POINTER(C.c_double), C.c_double, C.POINTER(C.c_double)]
    L.nbody_step_undo.argtypes = [vp]
    L.nbody_sync.argtypes = [vp]
    L.nbody_snapshot_async.argtypes = [vp, C.POINTER(C.c_uint64)]
    L.nbody_snapshot_wait.argtypes = [vp, C.c_uint64, C.POINTER(vp)]
    L.nbody_snapshot_release.argtypes = [vp, C.c_uint64]
    L.nbody_download_aos.argtypes = [vp, vp]
    L.nbody_energy.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.nbody_diagnostics.argtypes = [vp, C.POINTER(C.c_double)]
    L.nbody_get_timing.argtypes = [vp, C.POINTER(Timing)]
    L.nbody_get_trace.argtypes = [vp, C.POINTER(TraceEntry), sz, C.POINTER(sz)]
    L.nbody_host_alloc.restype = vp
    L.nbody_host_alloc.argtypes = [sz]
    L.nbody_host_free.argtypes = [vp]
    L.nbody_host_free.restype = None
    L.nbody_measure_fma_peak.argtypes = [C.c_int, C.c_int, C.c_double, C.POINTER(C.c_double)]
    L.nbody_sym_variant_for.argtypes = [sz, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int),
                                        C.POINTER(sz), C.POINTER(sz)]
    if hasattr(L, "nbody_rhs_noalloc"):
        L.nbody_rhs_noalloc.argtypes = [C.c_int, C.c_int, sz, vp, vp, vp, C.c_double, C.c_double, vp]
    return L


def declared_symbols() -> list[str]:
    """Every function name include/nbody_b200.h declares."""
    import re
    txt = open(HEADER_PATH).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(nbody_[a-z0-9_]+)\s*\(", txt)))


def partition(n: int, world: int, rank: int):
    """(begin, end, block_rows) of rank's owned rows; ValueError where the reference throws."""
    b, e, blk = C.c_size_t(), C.c_size_t(), C.c_size_t()
    if lib().nbody_partition(n, world, rank, C.byref(b), C.byref(e), C.byref(blk)):
        raise ValueError(lib().nbody_last_error(None).decode())
    return b.value, e.value, blk.value


def sym_variant_for(n: int, precision: int = 64, world: int = 1, uniform_mass: bool = True,
                    deterministic: bool = False):
    """(variant, kernel_block_rows, rank_block_rows) of the automatic pair-symmetric choice."""
    v, kb, rb = C.c_int(), C.c_size_t(), C.c_size_t()
    if lib().nbody_sym_variant_for(n, precision, world, int(uniform_mass), int(deterministic),
                                   C.byref(v), C.byref(kb), C.byref(rb)):
        raise ValueError(lib().nbody_last_error(None).decode())
    return v.value, kb.value, rb.value


def nccl_unique_id() -> bytes:
    buf = C.create_string_buffer(NCCL_UID_BYTES)
    rc = lib().nbody_nccl_unique_id(buf)
    if rc:
        raise NBodyError(lib().nbody_last_error(None).decode())
    return buf.raw


NOALLOC_MAX_BODIES = 65536


def rhs_noalloc(n: int, m_dev: int, r4_dev: int, acc_dev: int, precision: int = 64, device: int = 0,
                G: float = 6.67e-11, eps: float = 1e-6, stream: int | None = None) -> None:
    """nbody_rhs_noalloc: accelerations of n bodies the caller already holds on the device.

    For hosts that cannot let the library reserve device memory (everything else here goes through
    a Context, whose creation calls cudaMalloc): `m_dev`, `r4_dev`, `acc_dev` are DEVICE addresses
    (ints) of T[n] masses, T4[n] positions and T[3n] output.  The work is enqueued on `stream`
    (a cudaStream_t as int; None = default stream) and NOT waited for.  n is limited to
    NOALLOC_MAX_BODIES because the kernel's working memory is a fixed array inside the library;
    calls that can run on the GPU at the same time share that array and must be ordered."""
    rc = lib().nbody_rhs_noalloc(device, precision, n, m_dev, r4_dev, acc_dev, G, eps, stream)
    if rc:
        raise NBodyError(lib().nbody_last_error(None).decode())


def measure_fma_peak(precision: int = 64, ms_target: float = 200.0, device: int = 0) -> float:
    out = C.c_double()
    rc = lib().nbody_measure_fma_peak(device, precision, ms_target, C.byref(out))
    if rc:
        raise NBodyError(lib().nbody_last_error(None).decode())
    return out.value


class PinnedArray:
    """numpy view over cudaHostAlloc memory (nbody_host_alloc)."""

    def __init__(self, shape, dtype):
        self.nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        self.ptr = lib().nbody_host_alloc(self.nbytes)
        if not self.ptr:
            raise NBodyError("nbody_host_alloc failed")
        buf = (C.c_char * self.nbytes).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=dtype).reshape(shape)

    def free(self):
        if self.ptr:
            self.array = None
            lib().nbody_host_free(self.ptr)
            self.ptr = None


class Context:
    """One GPU / one rank.  Thin RAII wrapper; every method is one C-ABI call."""

    def __init__(self, n: int, precision: int = 64, device: int = 0, rank: int = 0,
                 world: int = 1, nccl_uid: bytes | None = None):
        self.L = lib()
        self.n, self.precision = int(n), int(precision)
        self.dtype = np.float64 if precision == 64 else np.float32
        self.rank, self.world = rank, world
        h = C.c_void_p()
        uid = C.create_string_buffer(nccl_uid, NCCL_UID_BYTES) if nccl_uid else None
        rc = self.L.nbody_create(C.byref(h), n, precision, device, rank, world, uid)
        if rc:
            raise NBodyError(self.L.nbody_last_error(None).decode())
        self.h = h

    def _ck(self, rc):
        if rc:
            raise NBodyError(self.L.nbody_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            self.L.nbody_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def p2p_export(self) -> bytes:
        buf = C.create_string_buffer(P2P_HANDLE_BYTES)
        self._ck(self.L.nbody_p2p_export(self.h, buf))
        return buf.raw

    def p2p_connect(self, blobs: list[bytes]):
        """blobs: every rank's p2p_export() result, in rank order."""
        assert len(blobs) == self.world and all(len(b) == P2P_HANDLE_BYTES for b in blobs)
        self._ck(self.L.nbody_p2p_connect(self.h, b"".join(blobs)))

    def set_params(self, G=6.67e-11, eps=1e-6):
        self._ck(self.L.nbody_set_params(self.h, G, eps))

    def set_option(self, name: str, value: int):
        self._ck(self.L.nbody_set_option(self.h, name.encode(), value))

    def owned_range(self):
        b, e = C.c_size_t(), C.c_size_t()
        self._ck(self.L.nbody_owned_range(self.h, C.byref(b), C.byref(e)))
        return b.value, e.value

    def upload_aos(self, bodies):
        b = np.ascontiguousarray(bodies, dtype=self.dtype)
        assert b.shape == (self.n, 7), b.shape
        self._ck(self.L.nbody_upload_aos(self.h, b.ctypes.data))

    def upload_aos_owned(self, owned):
        """Collective: this rank's rows [owned_range()) only; the rest arrives by all-gather."""
        b, e = self.owned_range()
        o = np.ascontiguousarray(owned, dtype=self.dtype)
        assert o.shape == (e - b, 7), (o.shape, b, e)
        self._ck(self.L.nbody_upload_aos_owned(self.h, o.ctypes.data if o.size else None))

    def download_aos_owned(self, out=None):
        b, e = self.owned_range()
        if out is None:
            out = np.empty((e - b, 7), dtype=self.dtype)
        assert out.shape == (e - b, 7) and out.flags.c_contiguous
        self._ck(self.L.nbody_download_aos_owned(self.h, out.ctypes.data if out.size else None))
        return out

    def upload_soa(self, m, r4, v4):
        m = np.ascontiguousarray(m, dtype=self.dtype)
        r4 = np.ascontiguousarray(r4, dtype=self.dtype)
        v4 = np.ascontiguousarray(v4, dtype=self.dtype)
        assert m.shape == (self.n,) and r4.shape == (self.n, 4) and v4.shape == (self.n, 4)
        self._ck(self.L.nbody_upload_soa(self.h, m.ctypes.data, r4.ctypes.data, v4.ctypes.data))

    def rhs(self):
        """Accelerations (n,3); rows outside owned_range() are NaN."""
        out = np.full((self.n, 3), np.nan, dtype=self.dtype)
        self._ck(self.L.nbody_rhs(self.h, out.ctypes.data))
        return out

    def step(self, scheme: int, tau: float, n_steps: int = 1, sync: bool = False):
        self._ck(self.L.nbody_step(self.h, scheme, tau, n_steps))
        if sync:
            self.sync()

    def step_tableau(self, a_sub, b, scale: float, tau: float, n_steps: int = 1, sync: bool = False):
        """Explicit RK with a sub-diagonal tableau: k_{s+1} = f(y + tau a_sub[s] k_s),
        y += tau scale sum b[s] k_s."""
        a = (C.c_double * max(len(a_sub), 1))(*a_sub)
        bb = (C.c_double * len(b))(*b)
        assert len(b) == len(a_sub) + 1
        self._ck(self.L.nbody_step_tableau(self.h, len(b), a, bb, scale, tau, n_steps))
        if sync:
            self.sync()

    @staticmethod
    def _tableau(a, b, b_err=None):
        b = np.ascontiguousarray(b, dtype=np.float64)
        a = np.ascontiguousarray(a, dtype=np.float64)
        assert a.shape == (b.size, b.size), (a.shape, b.size)
        dp = C.POINTER(C.c_double)
        be = None if b_err is None else np.ascontiguousarray(b_err, dtype=np.float64)
        assert be is None or be.shape == b.shape
        return a, b, be, dp

    def step_rk(self, a, b, tau: float, n_steps: int = 1, sync: bool = False):
        """General explicit RK, whole tableau as data: a (s, s) strictly lower, b (s,)."""
        a, b, _, dp = self._tableau(a, b)
        self._ck(self.L.nbody_step_rk(self.h, b.size, a.ctypes.data_as(dp), b.ctypes.data_as(dp),
                                      tau, n_steps))
        if sync:
            self.sync()

    def step_embedded(self, a, b, b_err, tau: float) -> np.ndarray:
        """One step of an embedded pair; returns [max|err_r|, max|err_v|, max|r|, max|v|]."""
        a, b, be, dp = self._tableau(a, b, b_err)
        out = np.zeros(4)
        self._ck(self.L.nbody_step_embedded(self.h, b.size, a.ctypes.data_as(dp),
                                            b.ctypes.data_as(dp), be.ctypes.data_as(dp), tau,
                                            out.ctypes.data_as(dp)))
        return out

    def step_undo(self):
        self._ck(self.L.nbody_step_undo(self.h))

    def sync(self):
        self._ck(self.L.nbody_sync(self.h))

    def snapshot_async(self) -> int:
        t = C.c_uint64()
        self._ck(self.L.nbody_snapshot_async(self.h, C.byref(t)))
        return t.value

    def snapshot_wait(self, ticket: int, copy: bool = True):
        p = C.c_void_p()
        self._ck(self.L.nbody_snapshot_wait(self.h, ticket, C.byref(p)))
        nbytes = self.n * 3 * np.dtype(self.dtype).itemsize
        a = np.frombuffer((C.c_char * nbytes).from_address(p.value), dtype=self.dtype)
        a = a.reshape(self.n, 3)
        return a.copy() if copy else a

    def snapshot_release(self, ticket: int):
        self._ck(self.L.nbody_snapshot_release(self.h, ticket))

    def snapshot(self):
        t = self.snapshot_async()
        a = self.snapshot_wait(t)
        self.snapshot_release(t)
        return a

    def download_aos(self, out=None):
        if out is None:
            out = np.empty((self.n, 7), dtype=self.dtype)
        self._ck(self.L.nbody_download_aos(self.h, out.ctypes.data))
        return out

    def energy(self):
        k, u = C.c_double(), C.c_double()
        self._ck(self.L.nbody_energy(self.h, C.byref(k), C.byref(u)))
        return k.value, u.value

    def diagnostics(self) -> dict:
        d = (C.c_double * 9)()
        self._ck(self.L.nbody_diagnostics(self.h, d))
        v = list(d)
        return {"kinetic": v[0], "potential": v[1], "momentum": np.array(v[2:5]),
                "mass_moment": np.array(v[5:8]), "mass": v[8],
                "com": np.array(v[5:8]) / v[8] if v[8] else np.zeros(3)}

    def trace(self, cap: int = 4096) -> list[dict]:
        """Operations of the last call (option "trace" = 1): kind, stream, start_ms, dur_ms."""
        buf = (TraceEntry * cap)()
        n = C.c_size_t()
        self._ck(self.L.nbody_get_trace(self.h, buf, cap, C.byref(n)))
        return [{"kind": TRACE_KINDS[buf[k].kind], "stream": "compute" if buf[k].stream == 0 else "comm",
                 "start_ms": buf[k].start_ms, "dur_ms": buf[k].dur_ms} for k in range(min(n.value, cap))]

    def timing(self) -> Timing:
        t = Timing()
        self._ck(self.L.nbody_get_timing(self.h, C.byref(t)))
        return t

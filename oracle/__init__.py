"""ctypes bindings for the parity oracle (TEST INFRASTRUCTURE ONLY).

Two libraries, both built by ``oracle/Makefile``:

* ``oracle/libchain_oracle.so``  -- our CPU restatement (``chain_oracle.c``), instrumented (f/p/v, visited pairs).
* ``oracle/_ref/libmm2ref_chain.so`` -- the reference's own ``chain.c``/``kalloc.c``/``misc.c``/``kthread.c``
  compiled unmodified (``ref_harness.c`` only makes the entry points callable on caller-owned arrays).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference`` legs may
import this package.  Nothing under ``minimap2_chaindp_b200/`` does.

Anchors are ``numpy.uint64`` arrays of shape ``(n, 2)``: column 0 = ``mm128_t.x``, column 1 = ``mm128_t.y``
(minimap.h:48).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "libchain_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libmm2ref_chain.so")

SEED_DTYPE = np.dtype([("x", "<u8"), ("y", "<u8"), ("p", "<i4"), ("f", "<i4")])  # struct new_seed, minimap.h:51-55
assert SEED_DTYPE.itemsize == 24


class Params(C.Structure):
    """Argument list of mm_chain_dp (chain.c:23) + optional max_iter (0 = unbounded = reference behaviour)."""

    _fields_ = [(k, C.c_int32) for k in
                ("max_dist_x", "max_dist_y", "bw", "max_skip", "min_cnt", "min_sc", "is_cdna", "n_segs", "max_iter")]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


# DP parameters per preset in this fork (options.c:22-55,79-142, map.c:358-366)
PRESETS = {
    "map-ont": dict(max_dist_x=5000, max_dist_y=5000, bw=500, max_skip=25, min_cnt=3, min_sc=40, is_cdna=0, n_segs=1),
    "map-pb": dict(max_dist_x=5000, max_dist_y=5000, bw=500, max_skip=25, min_cnt=3, min_sc=40, is_cdna=0, n_segs=1),
    "asm5": dict(max_dist_x=5000, max_dist_y=5000, bw=500, max_skip=25, min_cnt=3, min_sc=40, is_cdna=0, n_segs=1),
    "ava-ont": dict(max_dist_x=10000, max_dist_y=10000, bw=500, max_skip=25, min_cnt=3, min_sc=100, is_cdna=0, n_segs=1),
    "ava-pb": dict(max_dist_x=10000, max_dist_y=10000, bw=2000, max_skip=25, min_cnt=3, min_sc=100, is_cdna=0, n_segs=1),
    "sr": dict(max_dist_x=500, max_dist_y=100, bw=100, max_skip=25, min_cnt=2, min_sc=25, is_cdna=0, n_segs=2),
    "splice": dict(max_dist_x=200000, max_dist_y=2000, bw=200000, max_skip=25, min_cnt=3, min_sc=40, is_cdna=1, n_segs=1),
}


def params(preset: str = "map-ont", **over) -> Params:
    d = dict(PRESETS[preset], max_iter=0)
    d.update(over)
    return Params(**d)


def build(quiet: bool = True) -> None:
    """Compile the restatement and (when /root/reference is present) the reference library."""
    subprocess.run(["make", "-C", HERE, "all"], check=True,
                   stdout=subprocess.DEVNULL if quiet else None, stderr=subprocess.DEVNULL if quiet else None)


_P = C.POINTER
_vp, _i64, _i32 = C.c_void_p, C.c_int64, C.c_int32
_orc = None
_ref = None


def _anchors(a) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.uint64)
    if a.ndim != 2 or a.shape[1] != 2:
        a = a.reshape(-1, 2)
    return a


def _ptr(arr) -> C.c_void_p:
    return C.c_void_p(arr.ctypes.data)


def _take(ptr, n, dtype, shape_tail=(), free=None):
    """Copy n items out of a callee-malloc()ed array and free it."""
    if not ptr or n == 0:
        if ptr and free is not None:
            free(ptr)
        return np.zeros((0,) + shape_tail, dtype=dtype)
    count = int(n) * int(np.prod(shape_tail, dtype=np.int64)) if shape_tail else int(n)
    buf = (C.c_char * (count * np.dtype(dtype).itemsize)).from_address(ptr)
    out = np.frombuffer(buf, dtype=dtype, count=count).copy().reshape((int(n),) + shape_tail)
    if free is not None:
        free(ptr)
    return out


def oracle_lib():
    global _orc
    if _orc is None:
        if not os.path.exists(ORACLE_SO):
            build()
        L = C.CDLL(ORACLE_SO)
        L.orc_avg_qspan.restype = C.c_float
        L.orc_avg_qspan.argtypes = [_i64, _vp]
        L.orc_fill.restype = _i64
        L.orc_fill.argtypes = [_P(Params), _i64, _vp, _vp, _vp, _vp, _vp, _vp]
        L.orc_compact.restype = C.c_uint32
        L.orc_compact.argtypes = [_i32, _i64, _vp, _vp, _vp, _vp, _vp, _vp]
        L.orc_bottom.restype = _i32
        L.orc_bottom.argtypes = [_i32, _i32, C.c_uint32, _vp, _P(_vp), _P(_vp), _P(_i64)]
        L.orc_chain_dp.restype = _i32
        L.orc_chain_dp.argtypes = [_P(Params), _i64, _vp, _P(_vp), _P(_vp), _P(_i64)]
        L.orc_chain_batch.restype = _i64
        L.orc_chain_batch.argtypes = [_P(Params), _i64, _vp, _vp, _vp, _vp, _vp, _P(_vp), _P(_vp)]
        L.orc_count_visited.restype = _i64
        L.orc_count_visited.argtypes = [_P(Params), _i64, _vp, _vp]
        L.orc_radix_sort_64.argtypes = [_vp, _i64]
        L.orc_radix_sort_128x.argtypes = [_vp, _i64]
        L.orc_free.argtypes = [_vp]
        _orc = L
    return _orc


def have_ref() -> bool:
    return os.path.exists(REF_SO)


def ref_lib():
    """The compiled reference.  Raises FileNotFoundError if it was never built (and cannot be: no /root/reference)."""
    global _ref
    if _ref is None:
        if not os.path.exists(REF_SO):
            if os.path.isdir("/root/reference"):
                build()
            if not os.path.exists(REF_SO):
                raise FileNotFoundError(REF_SO)
        L = C.CDLL(REF_SO)
        L.ref_chain_dp.restype = _i32
        L.ref_chain_dp.argtypes = [_P(Params), _i64, _vp, _P(_vp), _P(_vp), _P(_i64)]
        L.ref_chain_dp_fpga.restype = C.c_uint32
        L.ref_chain_dp_fpga.argtypes = [_P(Params), _i64, _vp, _P(_vp)]
        L.ref_chain_dp_bottom.restype = _i32
        L.ref_chain_dp_bottom.argtypes = [_P(Params), C.c_uint32, _vp, _P(_vp), _P(_vp), _P(_i64)]
        L.ref_chain_batch_mt.restype = C.c_double
        L.ref_chain_batch_mt.argtypes = [_P(Params), _i64, _vp, _vp, C.c_int, C.c_int, _vp, _vp, _vp, _P(_vp), _P(_vp)]
        L.ref_free.argtypes = [_vp]
        L.radix_sort_64.argtypes = [_vp, _vp]
        L.radix_sort_128x.argtypes = [_vp, _vp]
        _ref = L
    return _ref


# ---------------------------------------------------------------------------------------------------------
# restatement (oracle proper)
# ---------------------------------------------------------------------------------------------------------

@dataclass
class Fill:
    f: np.ndarray
    p: np.ndarray
    v: np.ndarray
    visited: np.ndarray
    window: np.ndarray
    total_visited: int


def fill(prm: Params, a) -> Fill:
    """chain.c:52-90 with instrumentation."""
    a = _anchors(a)
    n = a.shape[0]
    f, p, v, vis, win = (np.zeros(n, dtype=np.int32) for _ in range(5))
    tot = oracle_lib().orc_fill(C.byref(prm), n, _ptr(a), _ptr(f), _ptr(p), _ptr(v), _ptr(vis), _ptr(win))
    return Fill(f, p, v, vis, win, int(tot))


def compact(min_sc: int, a, f, p, v):
    """chain.c:93-122 -> (records[new_i] as SEED_DTYPE, newid[n])."""
    a = _anchors(a)
    n = a.shape[0]
    rec = np.zeros(max(n, 1), dtype=SEED_DTYPE)
    newid = np.zeros(max(n, 1), dtype=np.int32)
    m = oracle_lib().orc_compact(min_sc, n, _ptr(a), _ptr(np.ascontiguousarray(f, np.int32)),
                                 _ptr(np.ascontiguousarray(p, np.int32)), _ptr(np.ascontiguousarray(v, np.int32)),
                                 _ptr(rec), _ptr(newid))
    return rec[:m].copy(), newid[:n].copy()


def bottom(prm: Params, rec):
    """chain.c:127-214 / mm_chain_dp_bottom -> (u[n_u], b[n_b,2])."""
    rec = np.ascontiguousarray(rec, dtype=SEED_DTYPE)
    L = oracle_lib()
    u, b, nb = _vp(), _vp(), _i64()
    nu = L.orc_bottom(prm.min_cnt, prm.min_sc, rec.shape[0], _ptr(rec), C.byref(u), C.byref(b), C.byref(nb))
    return (_take(u.value, nu, np.uint64, free=L.orc_free), _take(b.value, nb.value, np.uint64, (2,), free=L.orc_free))


def chain_dp(prm: Params, a):
    """Whole of mm_chain_dp -> (u[n_u], b[n_b,2])."""
    a = _anchors(a)
    L = oracle_lib()
    u, b, nb = _vp(), _vp(), _i64()
    nu = L.orc_chain_dp(C.byref(prm), a.shape[0], _ptr(a), C.byref(u), C.byref(b), C.byref(nb))
    return (_take(u.value, nu, np.uint64, free=L.orc_free), _take(b.value, nb.value, np.uint64, (2,), free=L.orc_free))


@dataclass
class BatchResult:
    n_u: np.ndarray      # int32 [n_reads]
    u_off: np.ndarray    # int64 [n_reads+1]
    b_off: np.ndarray    # int64 [n_reads+1]
    u: np.ndarray        # uint64 [sum n_u]   score<<32 | cnt, chains in first-anchor-x order per read
    b: np.ndarray        # uint64 [sum cnt, 2]
    visited: int = 0     # metric numerator (0 when the producer does not count)
    seconds: float = 0.0

    def same_as(self, o: "BatchResult") -> bool:
        return (np.array_equal(self.n_u, o.n_u) and np.array_equal(self.u_off, o.u_off)
                and np.array_equal(self.b_off, o.b_off) and np.array_equal(self.u, o.u) and np.array_equal(self.b, o.b))


def chain_batch(prm: Params, offs, a) -> BatchResult:
    a = _anchors(a)
    offs = np.ascontiguousarray(offs, dtype=np.int64)
    nr = offs.shape[0] - 1
    n_u = np.zeros(nr, np.int32)
    u_off = np.zeros(nr + 1, np.int64)
    b_off = np.zeros(nr + 1, np.int64)
    L = oracle_lib()
    u, b = _vp(), _vp()
    vis = L.orc_chain_batch(C.byref(prm), nr, _ptr(offs), _ptr(a), _ptr(n_u), _ptr(u_off), _ptr(b_off), C.byref(u), C.byref(b))
    return BatchResult(n_u, u_off, b_off, _take(u.value, u_off[-1], np.uint64, free=L.orc_free),
                       _take(b.value, b_off[-1], np.uint64, (2,), free=L.orc_free), visited=int(vis))


def count_visited(prm: Params, offs, a) -> int:
    a = _anchors(a)
    offs = np.ascontiguousarray(offs, dtype=np.int64)
    return int(oracle_lib().orc_count_visited(C.byref(prm), offs.shape[0] - 1, _ptr(offs), _ptr(a)))


def radix_sort_64(z):
    z = np.ascontiguousarray(z, dtype=np.uint64).copy()
    oracle_lib().orc_radix_sort_64(_ptr(z), z.shape[0])
    return z


def radix_sort_128x(z):
    z = _anchors(z).copy()
    oracle_lib().orc_radix_sort_128x(_ptr(z), z.shape[0])
    return z


# ---------------------------------------------------------------------------------------------------------
# compiled reference
# ---------------------------------------------------------------------------------------------------------

def ref_chain_dp(prm: Params, a):
    a = _anchors(a)
    L = ref_lib()
    u, b, nb = _vp(), _vp(), _i64()
    nu = L.ref_chain_dp(C.byref(prm), a.shape[0], _ptr(a), C.byref(u), C.byref(b), C.byref(nb))
    return (_take(u.value, nu, np.uint64, free=L.ref_free), _take(b.value, nb.value, np.uint64, (2,), free=L.ref_free))


def ref_chain_dp_fpga(prm: Params, a):
    """mm_chain_dp_fpga (chain.c:218) -> records[new_i]."""
    a = _anchors(a)
    L = ref_lib()
    rec = _vp()
    m = L.ref_chain_dp_fpga(C.byref(prm), a.shape[0], _ptr(a), C.byref(rec))
    return _take(rec.value, m, SEED_DTYPE, free=L.ref_free)


def ref_chain_dp_bottom(prm: Params, rec):
    rec = np.ascontiguousarray(rec, dtype=SEED_DTYPE).copy()
    L = ref_lib()
    u, b, nb = _vp(), _vp(), _i64()
    nu = L.ref_chain_dp_bottom(C.byref(prm), rec.shape[0], _ptr(rec), C.byref(u), C.byref(b), C.byref(nb))
    return (_take(u.value, nu, np.uint64, free=L.ref_free), _take(b.value, nb.value, np.uint64, (2,), free=L.ref_free))


def ref_chain_batch(prm: Params, offs, a, n_threads: int = 1, use_kalloc: bool = True, keep: bool = True) -> BatchResult:
    """Reference mm_chain_dp once per read under the reference's kt_for; .seconds = kt_for wall time."""
    a = _anchors(a)
    offs = np.ascontiguousarray(offs, dtype=np.int64)
    nr = offs.shape[0] - 1
    n_u = np.zeros(nr, np.int32)
    u_off = np.zeros(nr + 1, np.int64)
    b_off = np.zeros(nr + 1, np.int64)
    L = ref_lib()
    u, b = _vp(), _vp()
    sec = L.ref_chain_batch_mt(C.byref(prm), nr, _ptr(offs), _ptr(a), int(n_threads), int(bool(use_kalloc)),
                               _ptr(n_u), _ptr(u_off), _ptr(b_off),
                               C.byref(u) if keep else None, C.byref(b) if keep else None)
    if keep:
        U = _take(u.value, u_off[-1], np.uint64, free=L.ref_free)
        B = _take(b.value, b_off[-1], np.uint64, (2,), free=L.ref_free)
    else:
        U, B = np.zeros(0, np.uint64), np.zeros((0, 2), np.uint64)
    return BatchResult(n_u, u_off, b_off, U, B, seconds=float(sec))


def ref_radix_sort_64(z):
    z = np.ascontiguousarray(z, dtype=np.uint64).copy()
    ref_lib().radix_sort_64(_ptr(z), C.c_void_p(z.ctypes.data + z.nbytes))
    return z


def ref_radix_sort_128x(z):
    z = _anchors(z).copy()
    ref_lib().radix_sort_128x(_ptr(z), C.c_void_p(z.ctypes.data + z.nbytes))
    return z

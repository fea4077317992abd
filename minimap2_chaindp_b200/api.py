"""Python mirror of the reference's chaining interface, on top of the C ABI (include/mm2chain_b200.h).

This is only a caller of ``libmm2chain_b200.so`` -- the ctypes stub a test or bench driver needs.  It contains no
chaining logic and there is no CPU fallback: if the CUDA library is missing or no B200 is visible every call raises.

Reference interface mirrored (file:line into the reference tree):

* ``mm_chain_dp``        chain.c:23  / mmpriv.h:68
* ``mm_chain_dp_fpga``   chain.c:218 / mmpriv.h:69
* ``mm_chain_dp_bottom`` chain.c:329 / mmpriv.h:70

Anchors are ``numpy.uint64`` arrays of shape ``(n, 2)`` (column 0 = ``mm128_t.x``, column 1 = ``mm128_t.y``).
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Optional

import numpy as np

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG, "libmm2chain_b200.so")

SEED_DTYPE = np.dtype([("x", "<u8"), ("y", "<u8"), ("p", "<i4"), ("f", "<i4")])  # struct new_seed (minimap.h:51-55)

OK, ERR_CUDA, ERR_ARG, ERR_NOMEM, ERR_STATE = 0, -1, -2, -3, -4

# every symbol include/mm2chain_b200.h declares (tests check the library exports all of them)
EXPORTED_SYMBOLS = (
    "mm_chain_dp", "mm_chain_dp_fpga", "mm_chain_dp_bottom", "mm2b200_entry_stats",
    "mm2b200_create", "mm2b200_destroy", "mm2b200_last_error", "mm2b200_device_count",
    "mm2b200_pinned_alloc", "mm2b200_pinned_free",
    "mm2b200_chain_batch", "mm2b200_fill_batch", "mm2b200_bottom_batch",
    "mm2b200_partition_reads", "mm2b200_chain_batch_multi",
    "mm2b200_chain_device", "mm2b200_fill_device", "mm2b200_launch_count", "mm2b200_stage_ms", "mm2b200_stage_reset", "mm2b200_int_peak",
    "mm2b200_set_split", "mm2b200_set_upload", "mm2b200_upload_stats", "mm2b200_wire_encode", "mm2b200_version",
)


class ChainError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"mm2chain_b200 error {code}: {msg}")
        self.code = code


class Params(C.Structure):
    """mm2b200_params_t: the argument list of mm_chain_dp (chain.c:23) + max_iter (0 = unbounded)."""

    _fields_ = [(k, C.c_int32) for k in
                ("max_dist_x", "max_dist_y", "bw", "max_skip", "min_cnt", "min_sc", "is_cdna", "n_segs", "max_iter")]


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


class _Config(C.Structure):
    _fields_ = [("device", C.c_int32), ("n_slots", C.c_int32), ("slot_anchors", C.c_int64),
                ("profile", C.c_int32), ("count_visited", C.c_int32)]


class _Result(C.Structure):
    _fields_ = [("n_reads", C.c_int64), ("n_u", C.c_void_p), ("u_off", C.c_void_p), ("b_off", C.c_void_p),
                ("u", C.c_void_p), ("b", C.c_void_p), ("n_units", C.c_int64), ("n_sub_batches", C.c_int64),
                ("n_visited", C.c_int64), ("n_lanes", C.c_int64)]


class _Records(C.Structure):
    _fields_ = [("n_reads", C.c_int64), ("rec_off", C.c_void_p), ("rec", C.c_void_p)]


class _DevResult(C.Structure):
    _fields_ = [("d_n_u", C.c_void_p), ("d_u_off", C.c_void_p), ("d_b_off", C.c_void_p),
                ("d_u", C.c_void_p), ("d_b", C.c_void_p), ("d_totals", C.c_void_p)]


_lib = None


def lib() -> C.CDLL:
    """Load libmm2chain_b200.so.  Raises (never falls back) when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(
                f"{LIB_PATH} is missing: build it with `python -m minimap2_chaindp_b200.build` "
                "(nvcc, sm_100a).  There is no CPU fallback for the chaining path.")
        # a failing drop-in symbol aborts the process by default (the reference has no error channel); under Python the
        # wrappers below raise ChainError from mm2b200_last_error() instead
        os.environ.setdefault("MM2B200_ON_ERROR", "null")
        L = C.CDLL(LIB_PATH)
        vp, i64, i32 = C.c_void_p, C.c_int64, C.c_int32
        L.mm2b200_entry_stats.argtypes = [C.POINTER(i64), C.POINTER(i64)]
        L.mm2b200_create.argtypes = [C.POINTER(_Config), C.POINTER(vp)]
        L.mm2b200_destroy.argtypes = [vp]
        L.mm2b200_destroy.restype = None
        L.mm2b200_last_error.restype = C.c_char_p
        L.mm2b200_version.restype = C.c_char_p
        L.mm2b200_device_count.restype = C.c_int
        L.mm2b200_pinned_alloc.restype = vp
        L.mm2b200_pinned_alloc.argtypes = [C.c_size_t]
        L.mm2b200_pinned_free.argtypes = [vp]
        L.mm2b200_pinned_free.restype = None
        L.mm2b200_chain_batch.argtypes = [vp, C.POINTER(Params), i64, vp, vp, C.POINTER(_Result)]
        L.mm2b200_fill_batch.argtypes = [vp, C.POINTER(Params), i64, vp, vp, C.POINTER(_Records), vp, vp, vp]
        L.mm2b200_bottom_batch.argtypes = [vp, C.POINTER(Params), i64, vp, vp, C.POINTER(_Result)]
        L.mm2b200_partition_reads.argtypes = [i64, vp, i32, vp]
        L.mm2b200_chain_batch_multi.argtypes = [vp, i32, C.POINTER(Params), i64, vp, vp, vp, vp]
        L.mm2b200_chain_device.argtypes = [vp, C.POINTER(Params), i64, i64, vp, vp, vp, C.POINTER(_DevResult)]
        L.mm2b200_launch_count.argtypes = [vp]
        L.mm2b200_launch_count.restype = i64
        L.mm2b200_stage_ms.argtypes = [vp, C.c_char_p, C.POINTER(C.c_double), C.POINTER(i64)]
        L.mm2b200_stage_reset.argtypes = [vp]
        L.mm2b200_stage_reset.restype = None
        L.mm2b200_int_peak.argtypes = [vp, C.c_int, C.POINTER(C.c_double)]
        L.mm2b200_set_upload.argtypes = [vp, i32, i32]
        L.mm2b200_set_split.argtypes = [vp, i32]
        L.mm2b200_upload_stats.argtypes = [vp, C.POINTER(i64), C.POINTER(i64)]
        L.mm2b200_wire_encode.argtypes = [vp, i64, vp, C.c_size_t] + [C.POINTER(C.c_size_t)] * 3
        # reference entry points
        L.mm_chain_dp.restype = vp
        L.mm_chain_dp.argtypes = [C.c_int] * 8 + [i64, vp, C.POINTER(C.c_int), C.POINTER(vp), vp]
        L.mm_chain_dp_fpga.restype = vp
        L.mm_chain_dp_fpga.argtypes = [C.c_int] * 7 + [i64, vp, C.POINTER(C.c_uint32)]
        L.mm_chain_dp_bottom.restype = vp
        L.mm_chain_dp_bottom.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(vp), vp, vp, C.c_uint32]
        _lib = L
    return _lib


def last_error() -> str:
    return (lib().mm2b200_last_error() or b"").decode()


def _check(rc: int) -> None:
    if rc != OK:
        raise ChainError(rc, last_error())


def _anchors(a) -> np.ndarray:
    a = np.asarray(a)
    if a.dtype != np.uint64 or not a.flags.c_contiguous:
        a = np.ascontiguousarray(a, dtype=np.uint64)
    return a.reshape(-1, 2)


def _view(ptr: Optional[int], count: int, dtype, tail=()) -> np.ndarray:
    """Copy ``count`` items out of context-owned memory."""
    dtype = np.dtype(dtype)
    per = int(np.prod(tail, dtype=np.int64)) if tail else 1
    if not ptr or count <= 0:
        return np.zeros((0,) + tuple(tail), dtype=dtype)
    buf = (C.c_char * (int(count) * per * dtype.itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype, count=int(count) * per).reshape((int(count),) + tuple(tail)).copy()


@dataclass
class BatchResult:
    """Same shape as oracle.BatchResult so parity tests can compare field by field."""
    n_u: np.ndarray
    u_off: np.ndarray
    b_off: np.ndarray
    u: np.ndarray
    b: np.ndarray
    n_units: int = 0
    n_sub_batches: int = 0
    n_visited: int = 0
    n_lanes: int = 0

    def read(self, r: int):
        return self.u[self.u_off[r]:self.u_off[r + 1]], self.b[self.b_off[r]:self.b_off[r + 1]]


class PinnedArray:
    """A numpy view over page-locked memory from mm2b200_pinned_alloc (freed with the object)."""

    def __init__(self, shape, dtype):
        self.dtype = np.dtype(dtype)
        self.shape = tuple(int(s) for s in (shape if isinstance(shape, (tuple, list)) else (shape,)))
        nbytes = max(int(np.prod(self.shape, dtype=np.int64)) * self.dtype.itemsize, 1)
        self.ptr = lib().mm2b200_pinned_alloc(nbytes)
        if not self.ptr:
            raise ChainError(ERR_NOMEM, last_error())
        self._buf = (C.c_char * nbytes).from_address(self.ptr)
        self.array = np.frombuffer(self._buf, dtype=self.dtype, count=int(np.prod(self.shape, dtype=np.int64))).reshape(self.shape)

    def free(self):
        if self.ptr:
            self.array = None
            self._buf = None
            lib().mm2b200_pinned_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class ChainContext:
    """mm2b200_ctx_t: one device, its streams, workspaces and pinned result arenas."""

    def __init__(self, device: int = 0, n_slots: int = 0, slot_anchors: int = 0, profile: bool = False,
                 count_visited: bool = False):
        self._h = C.c_void_p()
        cfg = _Config(device=device, n_slots=n_slots, slot_anchors=slot_anchors, profile=int(profile),
                      count_visited=int(count_visited))
        _check(lib().mm2b200_create(C.byref(cfg), C.byref(self._h)))
        self.device = device

    def close(self):
        if self._h:
            lib().mm2b200_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- Section 3 ------------------------------------------------------------------------------------------
    def chain_batch_raw(self, prm: Params, n_reads: int, offs_ptr: int, a_ptr: int) -> _Result:
        """mm2b200_chain_batch on raw host pointers; returns the ctypes result struct (context-owned memory)."""
        res = _Result()
        _check(lib().mm2b200_chain_batch(self._h, C.byref(prm), n_reads, offs_ptr, a_ptr, C.byref(res)))
        return res

    def chain_batch(self, prm: Params, offs, a) -> BatchResult:
        """mm_chain_dp (chain.c:23) over a batch of reads; read r owns a[offs[r]:offs[r+1]]."""
        a = _anchors(a)
        offs = np.ascontiguousarray(offs, dtype=np.int64)
        nr = offs.shape[0] - 1
        res = self.chain_batch_raw(prm, nr, offs.ctypes.data, a.ctypes.data)
        return self._result(res, nr)

    @staticmethod
    def _result(res: _Result, nr: int) -> BatchResult:
        u_off = _view(res.u_off, nr + 1, np.int64)
        b_off = _view(res.b_off, nr + 1, np.int64)
        if u_off.shape[0] == 0:
            u_off, b_off = np.zeros(1, np.int64), np.zeros(1, np.int64)
        return BatchResult(_view(res.n_u, nr, np.int32), u_off, b_off,
                           _view(res.u, int(u_off[-1]), np.uint64), _view(res.b, int(b_off[-1]), np.uint64, (2,)),
                           int(res.n_units), int(res.n_sub_batches), int(res.n_visited), int(res.n_lanes))

    def fill_batch(self, prm: Params, offs, a, want_fpv: bool = False):
        """mm_chain_dp_fpga (chain.c:218) over a batch -> (rec_off, records[, f, p, v])."""
        a = _anchors(a)
        offs = np.ascontiguousarray(offs, dtype=np.int64)
        nr = offs.shape[0] - 1
        n = int(offs[-1] - offs[0]) if nr > 0 else 0
        out = _Records()
        f = p = v = None
        if want_fpv:
            f, p, v = (np.zeros(max(n, 1), np.int32) for _ in range(3))
        _check(lib().mm2b200_fill_batch(self._h, C.byref(prm), nr, offs.ctypes.data, a.ctypes.data, C.byref(out),
                                         f.ctypes.data if want_fpv else None, p.ctypes.data if want_fpv else None,
                                         v.ctypes.data if want_fpv else None))
        rec_off = _view(out.rec_off, nr + 1, np.int64)
        if rec_off.shape[0] == 0:
            rec_off = np.zeros(1, np.int64)
        rec = _view(out.rec, int(rec_off[-1]), SEED_DTYPE)
        if want_fpv:
            return rec_off, rec, f[:n], p[:n], v[:n]
        return rec_off, rec

    def bottom_batch(self, prm: Params, rec_off, rec) -> BatchResult:
        """mm_chain_dp_bottom (chain.c:329) over a batch of record arrays."""
        rec = np.ascontiguousarray(rec, dtype=SEED_DTYPE)
        rec_off = np.ascontiguousarray(rec_off, dtype=np.int64)
        nr = rec_off.shape[0] - 1
        res = _Result()
        _check(lib().mm2b200_bottom_batch(self._h, C.byref(prm), nr, rec_off.ctypes.data, rec.ctypes.data, C.byref(res)))
        return self._result(res, nr)

    # -- Section 4 ------------------------------------------------------------------------------------------
    def chain_device(self, prm: Params, n_reads: int, n_anchors: int, d_offs_ptr: int, d_a_ptr: int, stream: int = 0) -> _DevResult:
        """Whole path on device-resident inputs, enqueued on ``stream`` (a cudaStream_t handle); no host sync."""
        out = _DevResult()
        _check(lib().mm2b200_chain_device(self._h, C.byref(prm), n_reads, n_anchors, d_offs_ptr, d_a_ptr,
                                           C.c_void_p(stream), C.byref(out)))
        return out

    def launch_count(self) -> int:
        return int(lib().mm2b200_launch_count(self._h))

    def stage_ms(self, stage: str):
        ms, n = C.c_double(), C.c_int64()
        _check(lib().mm2b200_stage_ms(self._h, stage.encode(), C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def stage_reset(self):
        lib().mm2b200_stage_reset(self._h)

    def set_split(self, on: bool = True):
        """chain_device: two halves of a batch on two streams (default) or every kernel alone on the caller's stream."""
        _check(lib().mm2b200_set_split(self._h, 1 if on else 0))

    def set_upload(self, wire: bool = True, threads: int = 0):
        """Host batcher: 8-byte wire format for the anchor upload (default on) and its encoder thread count (0 = all cores)."""
        _check(lib().mm2b200_set_upload(self._h, 1 if wire else 0, threads))

    def upload_stats(self):
        """(bytes copied host->device, sub-batches sent in the wire format) of the last batch call."""
        b, n = C.c_int64(), C.c_int64()
        _check(lib().mm2b200_upload_stats(self._h, C.byref(b), C.byref(n)))
        return b.value, n.value

    def int_peak(self, mode: int = 1) -> float:
        """Measured integer lane-ops/s of the device (mode 0: ALU pipe only; 1: ALU+FMA-pipe mix)."""
        v = C.c_double()
        _check(lib().mm2b200_int_peak(self._h, mode, C.byref(v)))
        return v.value


WIRE_BLOCK = 256


def wire_encode(a: np.ndarray):
    """mm2b200_wire_encode (host only): returns (packed u64[n], descriptors u64[blocks], dict u32[entries]) or None
    when the anchors do not fit the wire format."""
    a = _anchors(a)
    n = len(a)
    nblk = (n + WIRE_BLOCK - 1) // WIRE_BLOCK
    cap = n * 8 + nblk * 8 + nblk * 2 * WIRE_BLOCK * 4 + 8
    buf = np.zeros(cap, dtype=np.uint8)
    nb, pt, pd = C.c_size_t(), C.c_size_t(), C.c_size_t()
    rc = lib().mm2b200_wire_encode(a.ctypes.data, n, buf.ctypes.data, cap, C.byref(nb), C.byref(pt), C.byref(pd))
    if rc == 1:
        return None
    _check(rc)
    packed = buf[:n * 8].view("<u8").copy()
    tab = buf[pt.value:pt.value + nblk * 8].view("<u8").copy()
    d = buf[pd.value:nb.value].view("<u4").copy()
    return packed, tab, d


def wire_decode(packed: np.ndarray, tab: np.ndarray, d: np.ndarray) -> np.ndarray:
    """numpy restatement of k_unpack_wire (tests)."""
    n = len(packed)
    out = np.zeros((n, 2), dtype=np.uint64)
    if n == 0:
        return out
    t = tab[np.arange(n) // WIRE_BLOCK]
    pos = (t & np.uint64(0xffffffff)).astype(np.int64)
    nhx = ((t >> np.uint64(32)) & np.uint64(0xffff)).astype(np.int64)
    qbits = np.uint64(24) - (t >> np.uint64(48))
    mid = (packed >> np.uint64(32)) & np.uint64(0xffffff)
    hx = d[pos + (packed >> np.uint64(56)).astype(np.int64)].astype(np.uint64)
    hy = d[pos + nhx + (mid >> qbits).astype(np.int64)].astype(np.uint64)
    out[:, 0] = (hx << np.uint64(32)) | (packed & np.uint64(0xffffffff))
    out[:, 1] = (hy << np.uint64(32)) | (mid & ((np.uint64(1) << qbits) - np.uint64(1)))
    return out


def partition_reads(offs, n_parts: int) -> np.ndarray:
    """mm2b200_partition_reads: contiguous read ranges of near-equal anchor count -> cuts[n_parts+1]."""
    offs = np.ascontiguousarray(offs, dtype=np.int64)
    cuts = np.zeros(n_parts + 1, dtype=np.int64)
    _check(lib().mm2b200_partition_reads(offs.shape[0] - 1, offs.ctypes.data, n_parts, cuts.ctypes.data))
    return cuts


def chain_batch_multi(ctxs, prm: Params, offs, a):
    """mm2b200_chain_batch_multi: one batch sharded over several contexts (one per GPU), results merged in read order."""
    a = _anchors(a)
    offs = np.ascontiguousarray(offs, dtype=np.int64)
    nr, n = offs.shape[0] - 1, len(ctxs)
    handles = (C.c_void_p * n)(*[c._h for c in ctxs])
    cuts = np.zeros(n + 1, dtype=np.int64)
    res = (_Result * n)()
    _check(lib().mm2b200_chain_batch_multi(handles, n, C.byref(prm), nr, offs.ctypes.data, a.ctypes.data,
                                           cuts.ctypes.data, res))
    parts = [ChainContext._result(res[k], int(cuts[k + 1] - cuts[k])) for k in range(n)]
    from .shard import merge_results
    return merge_results(parts), cuts


# ---------------------------------------------------------------------------------------------------------
# Section 1: the reference's own entry points, called exactly as map.c would (malloc()ed input that the callee
# consumes, malloc()ed outputs the caller frees)
# ---------------------------------------------------------------------------------------------------------
_libc = C.CDLL(None)
_libc.malloc.restype = C.c_void_p
_libc.malloc.argtypes = [C.c_size_t]
_libc.free.argtypes = [C.c_void_p]


def _malloc_copy(arr: np.ndarray) -> int:
    p = _libc.malloc(max(arr.nbytes, 1))
    if arr.nbytes:
        C.memmove(p, arr.ctypes.data, arr.nbytes)
    return p


def entry_stats():
    """(device batches, calls) served so far by the combining dispatcher behind the three drop-in symbols."""
    nb, nc = C.c_int64(0), C.c_int64(0)
    lib().mm2b200_entry_stats(C.byref(nb), C.byref(nc))
    return int(nb.value), int(nc.value)


def mm_chain_dp(prm: Params, a):
    """chain.c:23 through the drop-in symbol (km = NULL).  Returns (u[n_u], b[n_b, 2])."""
    a = _anchors(a)
    L = lib()
    n_u, u = C.c_int(0), C.c_void_p()
    b = L.mm_chain_dp(prm.max_dist_x, prm.max_dist_y, prm.bw, prm.max_skip, prm.min_cnt, prm.min_sc, prm.is_cdna,
                      prm.n_segs, a.shape[0], _malloc_copy(a), C.byref(n_u), C.byref(u), None)
    if not b and last_error():      # NULL is also the legitimate "no chain" answer; the error string tells them apart
        raise ChainError(ERR_CUDA, last_error())
    U = _view(u.value, n_u.value, np.uint64)
    nb = int((U & np.uint64(0xffffffff)).sum()) if n_u.value else 0
    B = _view(b, nb, np.uint64, (2,))
    _libc.free(u)
    _libc.free(b)
    return U, B


def mm_chain_dp_fpga(prm: Params, a) -> np.ndarray:
    """chain.c:218 through the drop-in symbol.  Returns records[new_i]."""
    a = _anchors(a)
    new_i = C.c_uint32(0)
    rec = lib().mm_chain_dp_fpga(prm.max_dist_x, prm.max_dist_y, prm.bw, prm.max_skip, prm.min_sc, prm.is_cdna,
                                 prm.n_segs, a.shape[0], _malloc_copy(a), C.byref(new_i))
    if not rec:
        raise ChainError(ERR_CUDA, last_error())
    out = _view(rec, new_i.value, SEED_DTYPE)
    _libc.free(rec)
    return out


def mm_chain_dp_bottom(prm: Params, rec):
    """chain.c:329 through the drop-in symbol.  Returns (u[n_u], b[n_b, 2])."""
    rec = np.ascontiguousarray(rec, dtype=SEED_DTYPE)
    n_u, u = C.c_int(0), C.c_void_p()
    b = lib().mm_chain_dp_bottom(prm.min_cnt, prm.min_sc, prm.n_segs, C.byref(n_u), C.byref(u), None,
                                 rec.ctypes.data, rec.shape[0])
    if not b and last_error():
        raise ChainError(ERR_CUDA, last_error())
    U = _view(u.value, n_u.value, np.uint64)
    nb = int((U & np.uint64(0xffffffff)).sum()) if n_u.value else 0
    B = _view(b, nb, np.uint64, (2,))
    _libc.free(u)
    _libc.free(b)
    return U, B

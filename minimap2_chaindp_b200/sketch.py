"""Python caller of the sketching stage (include/mm2sketch_b200.h): bases in, minimizers out, on the device.

Only a ctypes stub over ``libmm2chain_b200.so``; no logic, no CPU fallback.  Mirrors ``mm_sketch`` (sketch.c:77) as
``collect_minimizers`` calls it per read (map.c:99-111).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import api
from .api import ChainContext, _check, _view

# every symbol include/mm2sketch_b200.h declares
EXPORTED_SYMBOLS = ("mm2b200_sketch_batch", "mm2b200_sketch_device", "mm2b200_sketch_device_noalloc", "mm2b200_sketch_stage_ms")

# (w, k, is_hpc) per preset (options.c:79-142: mm_set_opt)
SKETCH_PRESETS = {"map-ont": (10, 15, 0), "ava-ont": (5, 15, 0), "map-pb": (10, 19, 1), "ava-pb": (5, 19, 1), "asm5": (19, 19, 0), "sr": (11, 21, 0)}


class SketchParams(C.Structure):
    """mm2b200_sketch_params_t"""
    _fields_ = [("w", C.c_int32), ("k", C.c_int32), ("is_hpc", C.c_int32)]


class _SketchResult(C.Structure):
    _fields_ = [("n_reads", C.c_int64), ("mv_off", C.c_void_p), ("mv", C.c_void_p)]


class _DevSketch(C.Structure):
    _fields_ = [("d_mv_off", C.c_void_p), ("d_mv", C.c_void_p), ("d_totals", C.c_void_p)]


_bound = False


def _lib():
    global _bound
    L = api.lib()
    if not _bound:
        vp, i64 = C.c_void_p, C.c_int64
        L.mm2b200_sketch_batch.argtypes = [vp, C.POINTER(SketchParams), i64, vp, vp, vp, C.POINTER(_SketchResult)]
        L.mm2b200_sketch_device.argtypes = [vp, C.POINTER(SketchParams), i64, i64, vp, vp, vp, vp, C.POINTER(_DevSketch), C.POINTER(i64)]
        L.mm2b200_sketch_stage_ms.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(i64)]
        L.mm2b200_sketch_device_noalloc.argtypes = [C.POINTER(SketchParams), i64, i64, vp, vp, vp, vp, vp, i64, vp, vp, C.POINTER(i64)]
        L.mm2b200_sketch_device_noalloc.restype = C.c_int
        _bound = True
    return L


def sketch_batch(ctx: ChainContext, seqs, w: int, k: int, is_hpc: bool = False, rid=None):
    """mm_sketch over a list of byte strings (or (seq_off, packed bytes)); returns (mv_off, mv[n, 2] uint64)."""
    if isinstance(seqs, tuple):
        off, raw = np.ascontiguousarray(seqs[0], np.int64), np.ascontiguousarray(seqs[1], np.uint8)
    else:
        off = np.concatenate([[0], np.cumsum([len(s) for s in seqs])]).astype(np.int64)
        raw = np.frombuffer(b"".join(seqs), np.uint8) if len(seqs) else np.zeros(0, np.uint8)
    n = off.shape[0] - 1
    rid_p = None
    if rid is not None:
        rid = np.ascontiguousarray(rid, np.uint32)
        assert rid.shape[0] == n
        rid_p = rid.ctypes.data
    prm = SketchParams(w, k, int(is_hpc))
    res = _SketchResult()
    _check(_lib().mm2b200_sketch_batch(ctx._h, C.byref(prm), n, off.ctypes.data, raw.ctypes.data if raw.size else None, rid_p, C.byref(res)))
    mv_off = _view(res.mv_off, n + 1, np.int64)
    return mv_off, _view(res.mv, int(mv_off[-1]) if mv_off.size else 0, np.uint64, (2,))


def sketch_stage_ms(ctx: ChainContext):
    ms, n = C.c_double(), C.c_int64()
    _check(_lib().mm2b200_sketch_stage_ms(ctx._h, C.byref(ms), C.byref(n)))
    return ms.value, n.value

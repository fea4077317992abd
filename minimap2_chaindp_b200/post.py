"""Python caller of the post-processing stage (include/mm2post_b200.h): chains -> mm_reg1_t regions on the device.

Only a ctypes stub over ``libmm2chain_b200.so``; it holds no logic and there is no CPU fallback.

Reference interface mirrored (file:line into the reference tree):

* ``mm_gen_regs``                              hit.c:52-88
* ``chain_post`` (``mm_set_parent``, ``mm_select_sub``, ``mm_join_long``)   map.c:238-247, hit.c:109-165,230-247,290-345
* ``mm_reg1_t``                                minimap.h:100-115  -> ``REG_DTYPE``
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import api
from .api import ChainContext, Params, _anchors, _check, _view

# mm_reg1_t, minimap.h:100-115 (80 bytes; p is NULL on this path)
REG_DTYPE = np.dtype([(k, "<i4") for k in ("id", "cnt", "rid", "score", "qs", "qe", "rs", "re", "parent", "subsc", "as", "mlen",
                                             "blen", "n_sub", "score0")] + [("bits", "<u4"), ("hash", "<u4"), ("div", "<f4"), ("p", "<u8")])
assert REG_DTYPE.itemsize == 80

# every symbol include/mm2post_b200.h declares
EXPORTED_SYMBOLS = ("mm2b200_chain_regs_batch", "mm2b200_post_device", "mm2b200_post_stage_ms")

# option values per preset as the reference's mm_set_opt leaves them (options.c:22-142); k is the preset's k-mer length
# (min_diff = 2k, map.c:242).  tests/test_post_oracle.py checks this table against the reference's own options.c.
POST_PRESETS = {
    "map-ont": dict(flag=0, mask_level=0.5, pri_ratio=0.8, best_n=5, k=15, max_join_long=20000, max_join_short=2000, min_join_flank_sc=1000, min_cnt=3),
    "map-pb": dict(flag=0, mask_level=0.5, pri_ratio=0.8, best_n=5, k=19, max_join_long=20000, max_join_short=2000, min_join_flank_sc=1000, min_cnt=3),
    "ava-ont": dict(flag=0x800000 | 0x400 | 0x2 | 0x1, mask_level=0.5, pri_ratio=0.0, best_n=5, k=15, max_join_long=20000, max_join_short=2000, min_join_flank_sc=1000, min_cnt=3),
    "asm5": dict(flag=0, mask_level=0.5, pri_ratio=0.8, best_n=50, k=19, max_join_long=20000, max_join_short=2000, min_join_flank_sc=1000, min_cnt=3),
}


class PostParams(C.Structure):
    """mm2b200_post_params_t"""
    _fields_ = [("flag", C.c_int32), ("mask_level", C.c_float), ("pri_ratio", C.c_float), ("best_n", C.c_int32), ("min_diff", C.c_int32),
                ("max_join_long", C.c_int32), ("max_join_short", C.c_int32), ("min_join_flank_sc", C.c_int32), ("min_cnt", C.c_int32)]


def post_params(preset: str = "map-ont", **over) -> PostParams:
    d = dict(POST_PRESETS[preset])
    d.update(over)
    k = d.pop("k")
    d.setdefault("min_diff", 2 * k)
    return PostParams(**d)


class _Regs(C.Structure):
    _fields_ = [("n_reads", C.c_int64), ("reg_off", C.c_void_p), ("regs", C.c_void_p), ("a_off", C.c_void_p), ("a", C.c_void_p),
                ("n_chains", C.c_int64)]


class _DevRegs(C.Structure):
    _fields_ = [("d_reg_off", C.c_void_p), ("d_regs", C.c_void_p), ("d_n_regs", C.c_void_p), ("d_totals", C.c_void_p)]


_bound = False


def _lib():
    global _bound
    L = api.lib()
    if not _bound:
        vp, i64 = C.c_void_p, C.c_int64
        L.mm2b200_chain_regs_batch.argtypes = [vp, C.POINTER(Params), C.POINTER(PostParams), i64, vp, vp, vp, vp, C.c_int, C.POINTER(_Regs)]
        L.mm2b200_post_device.argtypes = [vp, C.POINTER(PostParams), i64, C.POINTER(api._DevResult), i64, vp, vp, vp, C.POINTER(_DevRegs)]
        L.mm2b200_post_stage_ms.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(i64)]
        _bound = True
    return L


@dataclass
class RegBatch:
    reg_off: np.ndarray
    regs: np.ndarray
    a_off: np.ndarray | None
    a: np.ndarray | None
    n_chains: int


def chain_regs_batch(ctx: ChainContext, prm: Params, post: PostParams, offs, a, hash_, qlen, want_anchors: bool = False) -> RegBatch:
    """mm_chain_dp + mm_gen_regs + chain_post over a batch (map.c:837-866 without the base alignment)."""
    offs = np.ascontiguousarray(offs, np.int64)
    a = _anchors(a)
    n = offs.shape[0] - 1
    hash_ = np.ascontiguousarray(hash_, np.uint32)
    qlen = np.ascontiguousarray(qlen, np.int32)
    assert hash_.shape[0] == n and qlen.shape[0] == n
    res = _Regs()
    _check(_lib().mm2b200_chain_regs_batch(ctx._h, C.byref(prm), C.byref(post), n, offs.ctypes.data, a.ctypes.data, hash_.ctypes.data,
                                            qlen.ctypes.data, int(want_anchors), C.byref(res)))
    reg_off = _view(res.reg_off, n + 1, np.int64)
    regs = _view(res.regs, int(reg_off[-1]) if n >= 0 and reg_off.size else 0, REG_DTYPE)
    a_off = a_out = None
    if want_anchors:
        a_off = _view(res.a_off, n + 1, np.int64)
        a_out = _view(res.a, int(a_off[-1]) if a_off.size else 0, np.uint64, (2,))
    return RegBatch(reg_off, regs, a_off, a_out, int(res.n_chains))


def post_device(ctx: ChainContext, post: PostParams, n_reads: int, chains: "api._DevResult", n_chains_cap: int, d_hash_ptr: int,
                d_qlen_ptr: int, stream: int = 0) -> _DevRegs:
    """Post-processing of device-resident chains (the result struct of ChainContext.chain_device, or one pointing at the
    caller's own device arrays); enqueued on ``stream``, no host sync."""
    out = _DevRegs()
    _check(_lib().mm2b200_post_device(ctx._h, C.byref(post), n_reads, C.byref(chains), n_chains_cap, d_hash_ptr, d_qlen_ptr,
                                       C.c_void_p(stream), C.byref(out)))
    return out


def post_stage_ms(ctx: ChainContext):
    ms, n = C.c_double(), C.c_int64()
    _check(_lib().mm2b200_post_stage_ms(ctx._h, C.byref(ms), C.byref(n)))
    return ms.value, n.value

"""ctypes bindings of the post-processing / sketching checkers (TEST INFRASTRUCTURE ONLY).

* ``chain_post``      -- our restatement (``post_oracle.c``) of mm_gen_regs + chain_post (hit.c:52-345, map.c:238-247)
* ``ref_chain_post``  -- the reference's own hit.c compiled unmodified (``oracle/_ref/libmm2ref_post.so``, ``ref_post_harness.c``)
* ``ref_sketch``      -- the reference's own mm_sketch (sketch.c:77)
* ``ref_preset``      -- the reference's own option presets (options.c)
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import HERE, _anchors, build, oracle_lib

REF_POST_SO = os.path.join(HERE, "_ref", "libmm2ref_post.so")

# mm_reg1_t, minimap.h:100-115 (80 bytes; p is NULL on this path)
REG_DTYPE = np.dtype([(k, "<i4") for k in ("id", "cnt", "rid", "score", "qs", "qe", "rs", "re", "parent", "subsc", "as", "mlen",
                                             "blen", "n_sub", "score0")] + [("bits", "<u4"), ("hash", "<u4"), ("div", "<f4"), ("p", "<u8")])
assert REG_DTYPE.itemsize == 80


class PostParams(C.Structure):
    """the mm_mapopt_t fields chain_post reads (same layout as mm2b200_post_params_t)"""
    _fields_ = [("flag", C.c_int32), ("mask_level", C.c_float), ("pri_ratio", C.c_float), ("best_n", C.c_int32), ("min_diff", C.c_int32),
                ("max_join_long", C.c_int32), ("max_join_short", C.c_int32), ("min_join_flank_sc", C.c_int32), ("min_cnt", C.c_int32)]


class Preset(C.Structure):
    _fields_ = [(k, C.c_int32) for k in ("k", "w", "is_hpc", "bucket_bits", "flag", "bw", "max_gap", "max_gap_ref", "max_frag_len",
                                         "max_chain_skip", "min_cnt", "min_chain_score")] + \
               [("mask_level", C.c_float), ("pri_ratio", C.c_float)] + \
               [(k, C.c_int32) for k in ("best_n", "max_join_long", "max_join_short", "min_join_flank_sc")] + \
               [("mid_occ_frac", C.c_float), ("mid_occ", C.c_int32), ("max_occ", C.c_int32)]


_ref = None


def have_ref() -> bool:
    return os.path.exists(REF_POST_SO)


def ref_lib():
    global _ref
    if _ref is None:
        if not os.path.exists(REF_POST_SO):
            if os.path.isdir("/root/reference"):
                build()
            if not os.path.exists(REF_POST_SO):
                raise FileNotFoundError(REF_POST_SO)
        L = C.CDLL(REF_POST_SO)
        L.ref_chain_post.argtypes = [C.POINTER(PostParams), C.c_uint32, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_sketch.restype = C.c_int64
        L.ref_sketch.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_int, C.c_void_p]
        L.ref_preset.argtypes = [C.c_char_p, C.POINTER(Preset)]
        assert L.ref_sizeof_reg() == REG_DTYPE.itemsize
        _ref = L
    return _ref


def _post(fn, pp: PostParams, hash_: int, qlen: int, u, a):
    u = np.ascontiguousarray(u, np.uint64)
    a = np.array(_anchors(a), copy=True)                      # rewritten in place (mm_squeeze_a, LONG_JOIN marks)
    regs = np.zeros(max(u.size, 1), REG_DTYPE)
    n = fn(C.byref(pp), C.c_uint32(int(hash_) & 0xffffffff), int(qlen), int(u.size), u.ctypes.data, a.ctypes.data, regs.ctypes.data)
    regs = regs[:n].copy()
    n_a = int((regs["as"] + regs["cnt"]).max()) if n else 0   # anchors still referenced after the squeeze
    return regs, a, n_a


def chain_post(pp: PostParams, hash_: int, qlen: int, u, a):
    """(regs, a after the call, anchors still referenced) from the restatement"""
    L = oracle_lib()
    L.orc_chain_post.argtypes = [C.POINTER(PostParams), C.c_uint32, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    assert L.orc_sizeof_reg() == REG_DTYPE.itemsize
    return _post(L.orc_chain_post, pp, hash_, qlen, u, a)


def ref_chain_post(pp: PostParams, hash_: int, qlen: int, u, a):
    """the same from the compiled reference"""
    return _post(ref_lib().ref_chain_post, pp, hash_, qlen, u, a)


def ref_sketch(seq: bytes, w: int, k: int, rid: int = 0, is_hpc: bool = False) -> np.ndarray:
    out = np.zeros((len(seq) + 1, 2), np.uint64)
    n = ref_lib().ref_sketch(seq, len(seq), w, k, rid, int(is_hpc), out.ctypes.data)
    return out[:n].copy()


def sketch(seq: bytes, w: int, k: int, rid: int = 0, is_hpc: bool = False) -> np.ndarray:
    """mm_sketch from the restatement (sketch_oracle.c)"""
    L = oracle_lib()
    L.orc_sketch.restype = C.c_int64
    L.orc_sketch.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_int, C.c_void_p]
    out = np.zeros((len(seq) + 1, 2), np.uint64)
    n = L.orc_sketch(seq, len(seq), w, k, rid, int(is_hpc), out.ctypes.data)
    return out[:n].copy()


def ref_preset(name: str = "") -> Preset:
    p = Preset()
    if ref_lib().ref_preset(name.encode(), C.byref(p)) != 0:
        raise KeyError(name)
    return p


# option values per preset as the reference's own mm_set_opt leaves them (options.c:22-142); checked against ref_preset in
# tests/test_post_oracle.py so that the GPU box (no reference) can use them
POST_PRESETS = {
    "map-ont": dict(flag=0, mask_level=0.5, pri_ratio=0.8, best_n=5, k=15, max_join_long=20000, max_join_short=2000, min_join_flank_sc=1000, min_cnt=3),
    "map-pb": dict(flag=0, mask_level=0.5, pri_ratio=0.8, best_n=5, k=19, max_join_long=20000, max_join_short=2000, min_join_flank_sc=1000, min_cnt=3),
    "ava-ont": dict(flag=0x800000 | 0x400 | 0x2 | 0x1, mask_level=0.5, pri_ratio=0.0, best_n=5, k=15, max_join_long=20000, max_join_short=2000, min_join_flank_sc=1000, min_cnt=3),
    "asm5": dict(flag=0, mask_level=0.5, pri_ratio=0.8, best_n=50, k=19, max_join_long=20000, max_join_short=2000, min_join_flank_sc=1000, min_cnt=3),
}


def post_params(preset: str = "map-ont", **over) -> PostParams:
    d = dict(POST_PRESETS[preset])
    d.update(over)
    k = d.pop("k")
    d.setdefault("min_diff", 2 * k)                            # mi->k * 2, map.c:242
    return PostParams(**d)

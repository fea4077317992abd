"""Python caller of the seeding stage (include/mm2seed_b200.h): index streams in HBM, minimizers in, chains out.

Like ``api.py`` this is only a ctypes stub over ``libmm2chain_b200.so``; it holds no seeding logic and there is no CPU
fallback.

Reference interface mirrored (file:line into the reference tree):

* ``fpga_load_index(addr, size, type)``      fpga.h:62, main.c:201-204   -> ``SeedIndex.load``
* ``fpga_set_params(...)``                    fpga.h:58, main.c:243       -> ``SeedParams``
* ``collect_seed_hits`` + ``mm_chain_dp``     map.c:187-236, chain.c:23   -> ``SeedIndex.seed_chain_batch``
* ``fpga_work`` (request -> response packet)  map.c:484-568               -> ``SeedIndex.work``
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import api
from .api import BatchResult, ChainContext, Params, SEED_DTYPE, _check, _view

INDEX_B, INDEX_H, INDEX_V, INDEX_P = 4, 5, 6, 7

# every symbol include/mm2seed_b200.h declares
EXPORTED_SYMBOLS = (
    "mm2b200_index_create", "mm2b200_index_destroy", "mm2b200_index_load", "mm2b200_index_commit", "mm2b200_index_reset", "mm2b200_index_info",
    "mm2b200_seed_chain_batch", "mm2b200_seed_fill_batch", "mm2b200_seed_work",
    "mm2b200_seed_launch_count", "mm2b200_seed_stage_ms",
)
# include/mm2map_b200.h
MAP_EXPORTED_SYMBOLS = ("mm2b200_map_regs_batch",)


class SeedParams(C.Structure):
    """mm2b200_seed_params_t"""
    _fields_ = [("chain", Params), ("flag", C.c_int32), ("max_occ", C.c_int32)]


class _SeedResult(C.Structure):
    _fields_ = [("chains", api._Result), ("n_anchors", C.c_int64), ("rep_len", C.c_void_p), ("mini_off", C.c_void_p),
                ("mini_pos", C.c_void_p), ("a_off", C.c_void_p), ("a", C.c_void_p), ("n_exact_sorts", C.c_int64)]


class _SeedRecords(C.Structure):
    _fields_ = [("n_anchors", C.c_int64), ("rec_off", C.c_void_p), ("rec", C.c_void_p), ("rep_len", C.c_void_p),
                ("mini_off", C.c_void_p), ("mini_pos", C.c_void_p)]


class _MapResult(C.Structure):
    _fields_ = [("n_reads", C.c_int64), ("reg_off", C.c_void_p), ("regs", C.c_void_p), ("rep_len", C.c_void_p),
                ("n_minimizers", C.c_int64), ("n_anchors", C.c_int64), ("n_chains", C.c_int64)]


@dataclass
class MapBatch:
    reg_off: np.ndarray
    regs: np.ndarray
    rep_len: np.ndarray
    n_minimizers: int
    n_anchors: int
    n_chains: int


_bound = False


def _lib():
    global _bound
    L = api.lib()
    if not _bound:
        vp, i64, i32 = C.c_void_p, C.c_int64, C.c_int32
        L.mm2b200_index_create.argtypes = [vp, C.POINTER(vp)]
        L.mm2b200_index_destroy.argtypes = [vp]
        L.mm2b200_index_destroy.restype = None
        L.mm2b200_index_load.argtypes = [vp, vp, i64, C.c_int]
        L.mm2b200_index_commit.argtypes = [vp]
        L.mm2b200_index_reset.argtypes = [vp]
        L.mm2b200_index_info.argtypes = [vp, C.POINTER(i32), C.POINTER(i64)]
        L.mm2b200_seed_chain_batch.argtypes = [vp, C.POINTER(SeedParams), i64, vp, vp, vp, vp, C.c_int, C.POINTER(_SeedResult)]
        L.mm2b200_seed_fill_batch.argtypes = [vp, C.POINTER(SeedParams), i64, vp, vp, vp, vp, C.POINTER(_SeedRecords)]
        L.mm2b200_seed_work.argtypes = [vp, C.POINTER(SeedParams), vp, i64, vp, i64, C.POINTER(i64)]
        L.mm2b200_seed_launch_count.argtypes = [vp]
        L.mm2b200_seed_launch_count.restype = i64
        L.mm2b200_seed_stage_ms.argtypes = [vp] + [C.POINTER(C.c_double)] * 3
        L.mm2b200_map_regs_batch.argtypes = [vp, vp, C.POINTER(SeedParams), vp, i64, vp, vp, vp, vp, C.POINTER(_MapResult)]
        _bound = True
    return L


@dataclass
class SeedBatch:
    chains: BatchResult
    n_anchors: int
    rep_len: np.ndarray
    mini_off: np.ndarray
    mini_pos: np.ndarray
    a_off: np.ndarray
    a: np.ndarray | None
    n_exact_sorts: int


@dataclass
class SeedRecords:
    n_anchors: int
    rec_off: np.ndarray
    rec: np.ndarray
    rep_len: np.ndarray
    mini_off: np.ndarray
    mini_pos: np.ndarray


def _u8(x) -> np.ndarray:
    if isinstance(x, (bytes, bytearray, memoryview)):
        return np.frombuffer(bytes(x), np.uint8)
    return np.ascontiguousarray(x).view(np.uint8).reshape(-1)


class SeedIndex:
    """mm2b200_index_t bound to a ChainContext."""

    def __init__(self, ctx: ChainContext):
        self.ctx = ctx
        self._h = C.c_void_p()
        _check(_lib().mm2b200_index_create(ctx._h, C.byref(self._h)))

    def close(self):
        if self._h:
            _lib().mm2b200_index_destroy(self._h)
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

    def load(self, data, kind: int, chunk: int = 0) -> None:
        """fpga_load_index: append a stream (optionally in chunks of ``chunk`` bytes, as load_index_buf does)."""
        d = _u8(data)
        step = chunk if chunk > 0 else max(d.size, 1)
        for at in range(0, max(d.size, 1), step):
            part = d[at:at + step]
            _check(_lib().mm2b200_index_load(self._h, part.ctypes.data if part.size else None, part.size, kind))

    def commit(self) -> None:
        _check(_lib().mm2b200_index_commit(self._h))

    def load_streams(self, B, H, V, P) -> "SeedIndex":
        for d, k in ((B, INDEX_B), (H, INDEX_H), (V, INDEX_V), (P, INDEX_P)):
            self.load(d, k)
        self.commit()
        return self

    def info(self):
        b, n = C.c_int32(), C.c_int64()
        _check(_lib().mm2b200_index_info(self._h, C.byref(b), C.byref(n)))
        return int(b.value), int(n.value)

    @staticmethod
    def _inputs(mv_off, mv, bid, qlen):
        mv = np.ascontiguousarray(mv, dtype=np.uint64).reshape(-1, 2)
        mv_off = np.ascontiguousarray(mv_off, dtype=np.int64)
        bid = np.ascontiguousarray(np.asarray(bid, dtype=np.int64) & 0xffffffff, dtype=np.uint32)
        qlen = np.ascontiguousarray(qlen, dtype=np.int32)
        nr = mv_off.shape[0] - 1
        assert bid.shape[0] == nr and qlen.shape[0] == nr
        return mv_off, mv, bid, qlen, nr

    def seed_chain_batch(self, prm: SeedParams, mv_off, mv, bid, qlen, want_anchors: bool = False) -> SeedBatch:
        """collect_seed_hits (map.c:187-236) + mm_chain_dp (chain.c:23) for a batch of reads given as minimizers."""
        mv_off, mv, bid, qlen, nr = self._inputs(mv_off, mv, bid, qlen)
        res = _SeedResult()
        _check(_lib().mm2b200_seed_chain_batch(self._h, C.byref(prm), nr, mv_off.ctypes.data, mv.ctypes.data, bid.ctypes.data,
                                               qlen.ctypes.data, int(want_anchors), C.byref(res)))
        chains = ChainContext._result(res.chains, nr)
        mini_off = _view(res.mini_off, nr + 1, np.int64)
        a_off = _view(res.a_off, nr + 1, np.int64)
        return SeedBatch(chains=chains, n_anchors=int(res.n_anchors), rep_len=_view(res.rep_len, nr, np.int32), mini_off=mini_off,
                         mini_pos=_view(res.mini_pos, int(mini_off[-1]) if nr >= 0 else 0, np.uint64), a_off=a_off,
                         a=_view(res.a, int(res.n_anchors), np.uint64, (2,)) if want_anchors else None,
                         n_exact_sorts=int(res.n_exact_sorts))

    def seed_fill_batch(self, prm: SeedParams, mv_off, mv, bid, qlen) -> SeedRecords:
        """collect_seed_hits + mm_chain_dp_fpga (chain.c:218): what the fork's device sends back, per read."""
        mv_off, mv, bid, qlen, nr = self._inputs(mv_off, mv, bid, qlen)
        res = _SeedRecords()
        _check(_lib().mm2b200_seed_fill_batch(self._h, C.byref(prm), nr, mv_off.ctypes.data, mv.ctypes.data, bid.ctypes.data,
                                              qlen.ctypes.data, C.byref(res)))
        rec_off = _view(res.rec_off, nr + 1, np.int64)
        mini_off = _view(res.mini_off, nr + 1, np.int64)
        return SeedRecords(n_anchors=int(res.n_anchors), rec_off=rec_off, rec=_view(res.rec, int(rec_off[-1]), SEED_DTYPE),
                           rep_len=_view(res.rep_len, nr, np.int32), mini_off=mini_off,
                           mini_pos=_view(res.mini_pos, int(mini_off[-1]), np.uint64))

    def work(self, prm: SeedParams, packet: bytes, cap: int = 16 << 20) -> bytes:
        """fpga_work (map.c:484-568): request packet in, response packet out."""
        req = np.frombuffer(packet, np.uint8)
        out = np.zeros(cap, np.uint8)
        n = C.c_int64()
        _check(_lib().mm2b200_seed_work(self._h, C.byref(prm), req.ctypes.data, req.size, out.ctypes.data, cap, C.byref(n)))
        return out[:n.value].tobytes()

    def map_regs_batch(self, sketch_prm, prm: SeedParams, post_prm, seq_off, seq, bid, hash_) -> MapBatch:
        """bases in, regions out (include/mm2map_b200.h): mm_sketch -> collect_seed_hits -> mm_chain_dp -> mm_gen_regs + chain_post.
        ``sketch_prm`` / ``post_prm`` are sketch.SketchParams / post.PostParams; ``raw_call`` of the same is used by bench.py."""
        from .post import REG_DTYPE
        seq_off = np.ascontiguousarray(seq_off, np.int64)
        seq = np.ascontiguousarray(seq, np.uint8)
        nr = seq_off.shape[0] - 1
        bid = np.ascontiguousarray(np.asarray(bid, dtype=np.int64) & 0xffffffff, dtype=np.uint32)
        hash_ = np.ascontiguousarray(np.asarray(hash_, dtype=np.int64) & 0xffffffff, dtype=np.uint32)
        assert bid.shape[0] == nr and hash_.shape[0] == nr
        res = _MapResult()
        _check(_lib().mm2b200_map_regs_batch(self._h, C.byref(sketch_prm), C.byref(prm), C.byref(post_prm), nr, seq_off.ctypes.data,
                                             seq.ctypes.data if seq.size else None, bid.ctypes.data, hash_.ctypes.data, C.byref(res)))
        reg_off = _view(res.reg_off, nr + 1, np.int64)
        return MapBatch(reg_off=reg_off, regs=_view(res.regs, int(reg_off[-1]) if reg_off.size else 0, REG_DTYPE),
                        rep_len=_view(res.rep_len, nr, np.int32), n_minimizers=int(res.n_minimizers), n_anchors=int(res.n_anchors),
                        n_chains=int(res.n_chains))

    def launch_count(self) -> int:
        return int(_lib().mm2b200_seed_launch_count(self._h))

    def stage_ms(self):
        v = [C.c_double() for _ in range(3)]
        _check(_lib().mm2b200_seed_stage_ms(self._h, *[C.byref(x) for x in v]))
        return {"lookup": v[0].value, "expand": v[1].value, "sort": v[2].value}


def seed_params(chain: Params, flag: int, max_occ: int) -> SeedParams:
    p = SeedParams()
    C.memmove(C.byref(p.chain), C.byref(chain), C.sizeof(Params))
    p.flag, p.max_occ = int(flag), int(max_occ)
    return p

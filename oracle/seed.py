"""ctypes bindings for the seeding oracle (seed_oracle.c) and a builder of synthetic index streams.

TEST INFRASTRUCTURE ONLY, like the rest of ``oracle/``.

``SeedIndex`` wraps the four streams the fork ships to its device (B/H/V/P, index.c:603-720).  ``build_streams`` makes
such streams from arbitrary (minimizer -> positions) data so that tests can reach cases the small golden fixtures do not
hold: long position lists, long probe sequences, deleted slots, x ties in reads of more than 64 anchors.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _anchors, _ptr, oracle_lib

_vp, _i64, _i32 = C.c_void_p, C.c_int64, C.c_int32


class _View(C.Structure):
    _fields_ = [("B", _vp), ("H", _vp), ("V", _vp), ("P", _vp), ("nB", _i64), ("nH", _i64), ("nV", _i64), ("nP", _i64), ("b", C.c_int)]


def _lib():
    L = oracle_lib()
    if not getattr(L, "_seed_bound", False):
        L.orc_index_view.restype = C.c_int
        L.orc_index_view.argtypes = [C.POINTER(_View), _vp, _i64, _vp, _i64, _vp, _i64, _vp, _i64]
        L.orc_idx_get.restype = C.c_int
        L.orc_idx_get.argtypes = [C.POINTER(_View), C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(_i64), C.POINTER(C.c_int)]
        L.orc_collect_seed_hits.restype = _i64
        L.orc_collect_seed_hits.argtypes = [C.POINTER(_View), C.c_int, C.c_int, _vp, _i64, C.c_uint32, C.c_int, _vp,
                                            C.POINTER(_i32), C.POINTER(_i32), _vp, C.c_int]
        L.orc_seed_count.restype = _i64
        L.orc_seed_count.argtypes = [C.POINTER(_View), C.c_int, _vp, _i64]
        L._seed_bound = True
    return L


class SeedIndex:
    """The index as the device sees it: four byte streams."""

    def __init__(self, B, H, V, P):
        self.B, self.H, self.V, self.P = (np.ascontiguousarray(np.frombuffer(bytes(s), np.uint8) if not isinstance(s, np.ndarray) else s,
                                                               dtype=np.uint8) for s in (B, H, V, P))
        self.view = _View()
        rc = _lib().orc_index_view(C.byref(self.view), _ptr(self.B), self.B.size, _ptr(self.H), self.H.size,
                                   _ptr(self.V), self.V.size, _ptr(self.P), self.P.size)
        if rc != 0:
            raise ValueError("inconsistent index streams")

    @property
    def b(self) -> int:
        return int(self.view.b)

    def get(self, minier: int):
        """mm_idx_get: returns the list of packed positions of one minimizer."""
        val, at, single = C.c_uint64(), _i64(), C.c_int()
        n = _lib().orc_idx_get(C.byref(self.view), C.c_uint64(minier), C.byref(val), C.byref(at), C.byref(single))
        if single.value:
            return [int(val.value)]
        return [int(x) for x in self.P.view(np.uint64)[at.value:at.value + n]]


def collect_seed_hits(ix: SeedIndex, flag: int, max_occ: int, mv, bid: int, qlen: int, sort: bool = True):
    """collect_seed_hits (map.c:187-236) for one read.  Returns (anchors (n,2) u64, rep_len, mini_pos u64[])."""
    mv = _anchors(mv)
    L = _lib()
    cap = int(L.orc_seed_count(C.byref(ix.view), max_occ, _ptr(mv), mv.shape[0]))
    a = np.zeros((max(cap, 1), 2), np.uint64)
    mini = np.zeros(max(mv.shape[0], 1), np.uint64)
    rep, n_mini = _i32(), _i32()
    n = L.orc_collect_seed_hits(C.byref(ix.view), flag, max_occ, _ptr(mv), mv.shape[0], C.c_uint32(bid & 0xffffffff), qlen,
                                _ptr(a), C.byref(rep), C.byref(n_mini), _ptr(mini), int(sort))
    return a[:n].copy(), int(rep.value), mini[:n_mini.value].copy()


# ---------------------------------------------------------------------------------------------------------
# synthetic streams
# ---------------------------------------------------------------------------------------------------------

def pack_position(refid: int, refpos: int, strand: int, rank: int) -> int:
    """The fork's packed position: refid 21 | refpos 21 | strand 1 | rank 21 (map.c:208-214)."""
    return (refid & 0x1fffff) << 43 | (refpos & 0x1fffff) << 22 | (strand & 1) << 21 | (rank & 0x1fffff)


def build_streams(table: dict, b: int = 14, rng=None, load: float = 0.5, deleted: float = 0.0):
    """Index streams for ``table``: {minimizer value (<= 56 bits): [packed positions]}.

    Every bucket's hash table is laid out with khash's own probing from an empty table of the smallest power-of-two size
    that keeps the load under ``load`` (khash never exceeds 0.77), so mm_idx_get finds every key.  ``deleted`` > 0 marks
    that fraction of the free slots "deleted" (flag bit 0) with a junk key, which the probe must step over.
    """
    rng = rng or np.random.default_rng(0)
    nb = 1 << b
    buckets = [[] for _ in range(nb)]
    for m, pos in table.items():
        buckets[m & (nb - 1)].append((m >> b, list(pos)))
    Bs = np.zeros((nb, 2), np.uint64)
    H, V, P = [], [], []
    allh = allp = 0
    for i, ent in enumerate(buckets):
        if not ent:
            continue
        n_buckets = 4
        while len(ent) > load * n_buckets:
            n_buckets *= 2
        flags = np.full(n_buckets, 2, np.uint8)     # empty
        keys = np.zeros(n_buckets, np.uint64)
        vals = np.zeros(n_buckets, np.uint64)
        plist = []
        if deleted > 0:
            for s in range(n_buckets):
                if rng.random() < deleted:
                    flags[s] = 1
                    keys[s] = int(rng.integers(1 << 40))
        n_live = 0
        for hi, pos in ent:
            key = hi << 1
            s = hi & 0xffffffff & (n_buckets - 1)
            step = 0
            while flags[s] == 0:
                step += 1
                s = (s + step) & (n_buckets - 1)
            # khash's put reuses the first deleted slot on the probe path only when the key is absent further on;
            # taking the first non-live slot is what a fresh insert does
            flags[s] = 0
            n_live += 1
            if len(pos) == 1:
                keys[s] = key | 1
                vals[s] = pos[0]
            else:
                keys[s] = key
                vals[s] = len(plist) << 32 | len(pos)
                plist.extend(pos)
        # a deleted slot in front of a live key is fine; an empty one is not -- inserts above stop only at non-live
        # slots, so every key is reachable as long as slots before it on its path are live or deleted: make the
        # remaining non-live slots on any path "deleted" only if they were deleted before (already the case)
        tmp = (n_buckets + 7) // 8 * 8
        Bs[i, 1] = allh << 28 | allp >> 8
        Bs[i, 0] = (allp & 0xff) << 56 | n_buckets << 24
        for g in range(tmp // 8):
            word = 0
            base16 = (g * 8) & ~15
            for s in range(16):
                f = int(flags[base16 + s]) if base16 + s < n_buckets else 0
                word |= f << (2 * s)
            grp = bytearray(64)
            grp[0:4] = int(word).to_bytes(4, "little")
            for s in range(8):
                j = g * 8 + s
                if j < n_buckets:
                    grp[4 + 6 * s:10 + 6 * s] = (int(keys[j]) & ((1 << 48) - 1)).to_bytes(6, "little")
            H.append(bytes(grp))
            V.append(np.array([vals[g * 8 + s] if g * 8 + s < n_buckets else 0 for s in range(8)], np.uint64).tobytes())
        P.append(np.array(plist, np.uint64).tobytes())
        allh += tmp
        allp += len(plist)
    return SeedIndex(Bs.tobytes(), b"".join(H), b"".join(V), b"".join(P))

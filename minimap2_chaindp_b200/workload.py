"""Synthetic anchor workloads "G-anc" for the BASELINE.json configs.

ctypes wrapper over ``libmm2workload.so`` (csrc/workload.c).  One deterministic PRNG stream per read, so a rank
can generate exactly its own shard: ``generate(spec, first, n_reads)``.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, replace

import numpy as np

PKG = os.path.dirname(os.path.abspath(__file__))
WL_PATH = os.path.join(PKG, "libmm2workload.so")


class _Spec(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("read_len", C.c_int32), ("gap_max", C.c_int32), ("p_survive", C.c_double),
                ("p_drift", C.c_double), ("span_lo", C.c_int32), ("span_hi", C.c_int32), ("n_contigs", C.c_int32),
                ("contig_len", C.c_int32), ("p_background", C.c_double), ("bg_mean", C.c_double),
                ("p_repeat", C.c_double), ("repeat_radius", C.c_int32), ("repeat_band", C.c_int32),
                ("n_loci", C.c_int32), ("reserved", C.c_int32)]


@dataclass(frozen=True)
class WorkloadSpec:
    name: str
    preset: str            # DP parameter preset (api.PRESETS)
    n_reads: int
    seed: int
    read_len: int
    gap_max: int
    p_survive: float
    p_drift: float
    span_lo: int
    span_hi: int
    n_contigs: int
    contig_len: int
    p_background: float = 0.05
    bg_mean: float = 4.0
    p_repeat: float = 0.02
    repeat_radius: int = 2000
    repeat_band: int = 500
    n_loci: int = 1

    def c_spec(self) -> _Spec:
        return _Spec(self.seed, self.read_len, self.gap_max, self.p_survive, self.p_drift, self.span_lo, self.span_hi,
                     self.n_contigs, self.contig_len, self.p_background, self.bg_mean, self.p_repeat,
                     self.repeat_radius, self.repeat_band, self.n_loci, 0)

    def scaled(self, n_reads: int) -> "WorkloadSpec":
        return replace(self, n_reads=int(n_reads))


_SEED = 0x5EED0000

# BASELINE.json configs[1..4] (configs[0] is the MT-human/MT-orang fixture under tests/golden/)
CONFIGS = {
    # map-ont: 100k synthetic 10 kb Nanopore-like reads (~10 % error) vs 100 x 1 Mb contigs; k=15, w=10
    "map-ont-100k": WorkloadSpec("map-ont-100k", "map-ont", 100_000, _SEED + 2, 10_000, 10, 0.9 ** 15, 0.10 / 3,
                                 15, 15, 100, 1_000_000),
    # ava-ont: 50k x 10 kb reads, ~50 overlapping (target,strand) loci per query; k=15, w=5, both reads noisy
    "ava-ont-50k": WorkloadSpec("ava-ont-50k", "ava-ont", 50_000, _SEED + 3, 10_000, 5, 0.12, 0.20 / 3,
                                15, 15, 50_000, 10_000, p_background=0.01, n_loci=50),
    # map-pb ultra-long: 20k x 100 kb reads, HPC spans 19..40, dense (n >= 50k anchors per read)
    "map-pb-ul-20k": WorkloadSpec("map-pb-ul-20k", "map-pb", 20_000, _SEED + 4, 100_000, 3, 0.80, 0.15 / 3,
                                  19, 40, 100, 2_000_000, p_background=0.05, p_repeat=0.10),
    # asm5: 50 Mb contigs at 1 % divergence; k=19, w=19 -> one serial run of ~4 M anchors per contig
    "asm5-50mb": WorkloadSpec("asm5-50mb", "asm5", 8, _SEED + 5, 50_000_000, 19, 0.99 ** 19, 0.01 / 3,
                              19, 19, 8, 60_000_000, p_background=0.002, p_repeat=0.002),
}

_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(WL_PATH):
            from . import build
            build.build_workload()
        L = C.CDLL(WL_PATH)
        L.wl_count.restype = C.c_int64
        L.wl_count.argtypes = [C.POINTER(_Spec), C.c_int64, C.c_int64, C.c_void_p]
        L.wl_fill.restype = None
        L.wl_fill.argtypes = [C.POINTER(_Spec), C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def count(spec: WorkloadSpec, first: int = 0, n_reads: int | None = None) -> np.ndarray:
    """offs[n_reads+1] for reads [first, first+n_reads)."""
    n_reads = spec.n_reads - first if n_reads is None else n_reads
    offs = np.zeros(n_reads + 1, dtype=np.int64)
    cs = spec.c_spec()
    lib().wl_count(C.byref(cs), first, n_reads, offs.ctypes.data)
    return offs


def fill(spec: WorkloadSpec, offs: np.ndarray, out: np.ndarray, first: int = 0) -> None:
    """Write the anchors of reads [first, first+len(offs)-1) into out[(offs[-1], 2) uint64] (may be pinned)."""
    assert out.dtype == np.uint64 and out.flags.c_contiguous and out.size >= 2 * int(offs[-1])
    cs = spec.c_spec()
    lib().wl_fill(C.byref(cs), first, offs.shape[0] - 1, offs.ctypes.data, out.ctypes.data)


def generate(spec: WorkloadSpec, first: int = 0, n_reads: int | None = None):
    """-> (offs[n_reads+1] int64, anchors[(n, 2)] uint64)."""
    offs = count(spec, first, n_reads)
    a = np.zeros((int(offs[-1]), 2), dtype=np.uint64)
    fill(spec, offs, a, first)
    return offs, a


# ---------------------------------------------------------------------------------------------------------
# "G-idx": synthetic index streams + reads given as minimizers, for the seeding stage (csrc/workload_seed.c)
# ---------------------------------------------------------------------------------------------------------
class _IdxSpec(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("n_sites", C.c_int64), ("hash_bits", C.c_int32), ("b", C.c_int32), ("gap_max", C.c_int32),
                ("ref_len", C.c_int32), ("p_dup", C.c_double), ("p_hot", C.c_double), ("n_hot", C.c_int32), ("reserved", C.c_int32)]


class _ReadSpec(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("read_len", C.c_int32), ("span", C.c_int32), ("p_match", C.c_double), ("p_drift", C.c_double)]


@dataclass(frozen=True)
class SeedWorkloadSpec:
    name: str
    preset: str
    n_reads: int
    seed: int
    n_sites: int              # minimizer sites of the reference (~ bases * 2 / (w + 1))
    read_len: int
    hash_bits: int = 30       # 2k, k = 15
    b: int = 14               # mi->b
    gap_max: int = 10         # w
    ref_len: int = 2_000_000  # the fork packs refpos in 21 bits
    p_dup: float = 0.03
    p_hot: float = 0.004
    n_hot: int = 40
    span: int = 15
    p_match: float = 0.9 ** 15            # (1 - e)^k at 10 % error, as in G-anc map-ont-100k
    p_drift: float = 0.10 / 3
    max_occ: int = 50         # opt->mid_occ
    flag: int = 0


SEED_CONFIGS = {
    # 100k x 10 kb reads against ~100 Mb of reference (19 M sites): index streams ~ 0.6 GB, far beyond L2
    "map-ont-100k-seeded": SeedWorkloadSpec("map-ont-100k-seeded", "map-ont", 100_000, _SEED + 12, 19_000_000, 10_000),
    "map-ont-small-seeded": SeedWorkloadSpec("map-ont-small-seeded", "map-ont", 2_000, _SEED + 13, 400_000, 10_000),
}


class SeedWorkload:
    """The built index (streams as numpy views over C memory) and a sampler of reads."""

    def __init__(self, spec: SeedWorkloadSpec):
        L = lib()
        if not hasattr(L, "_seed_bound"):
            L.wls_build.restype = C.c_void_p
            L.wls_build.argtypes = [C.POINTER(_IdxSpec)]
            L.wls_free.argtypes = [C.c_void_p]
            L.wls_free.restype = None
            L.wls_streams.argtypes = [C.c_void_p] + [C.POINTER(C.c_void_p), C.POINTER(C.c_int64)] * 4 + [C.POINTER(C.c_int64)] * 2
            L.wls_streams.restype = None
            L.wls_reads_count.restype = C.c_int64
            L.wls_reads_count.argtypes = [C.c_void_p, C.POINTER(_ReadSpec), C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
            L.wls_reads_fill.restype = None
            L.wls_reads_fill.argtypes = [C.c_void_p, C.POINTER(_ReadSpec), C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]
            L._seed_bound = True
        self.spec = spec
        cs = _IdxSpec(spec.seed, spec.n_sites, spec.hash_bits, spec.b, spec.gap_max, spec.ref_len, spec.p_dup, spec.p_hot, spec.n_hot, 0)
        self._h = L.wls_build(C.byref(cs))
        if not self._h:
            raise MemoryError("wls_build failed")
        ptr = [C.c_void_p() for _ in range(4)]
        size = [C.c_int64() for _ in range(4)]
        nk, mo = C.c_int64(), C.c_int64()
        args = []
        for p, s in zip(ptr, size):
            args += [C.byref(p), C.byref(s)]
        L.wls_streams(self._h, *args, C.byref(nk), C.byref(mo))
        self.streams = []
        for p, s in zip(ptr, size):
            n = int(s.value)
            self.streams.append(np.frombuffer((C.c_char * max(n, 1)).from_address(p.value), np.uint8, n) if n else np.zeros(0, np.uint8))
        self.n_keys, self.max_occ_seen = int(nk.value), int(mo.value)
        self._rs = _ReadSpec(spec.seed ^ 0xABCDEF, spec.read_len, spec.span, spec.p_match, spec.p_drift)

    @property
    def B(self): return self.streams[0]
    @property
    def H(self): return self.streams[1]
    @property
    def V(self): return self.streams[2]
    @property
    def P(self): return self.streams[3]

    def reads(self, first: int = 0, n_reads: int | None = None, out: np.ndarray | None = None):
        """-> (mv_off[n+1] int64, mv[(m, 2)] uint64, bid[n] uint32, qlen[n] int32) for reads [first, first + n)."""
        n = self.spec.n_reads - first if n_reads is None else n_reads
        off = np.zeros(n + 1, np.int64)
        qlen = np.zeros(max(n, 1), np.int32)
        bid = np.zeros(max(n, 1), np.uint32)
        m = lib().wls_reads_count(self._h, C.byref(self._rs), first, n, off.ctypes.data, qlen.ctypes.data, bid.ctypes.data)
        mv = out if out is not None else np.zeros((int(m), 2), np.uint64)
        assert mv.dtype == np.uint64 and mv.size >= 2 * int(m)
        lib().wls_reads_fill(self._h, C.byref(self._rs), first, n, off.ctypes.data, mv.ctypes.data)
        return off, mv, bid[:n], qlen[:n]

    def close(self):
        if self._h:
            self.streams = []
            lib().wls_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---------------------------------------------------------------------------------------------------------
# "G-seq": a synthetic genome and reads as BASES, for the composed pipeline (csrc/workload_seed.c, wlg_*)
# ---------------------------------------------------------------------------------------------------------
class _GenomeSpec(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("n_contigs", C.c_int32), ("contig_len", C.c_int32), ("p_repeat", C.c_double),
                ("read_len", C.c_int32), ("err", C.c_double)]


@dataclass(frozen=True)
class SeqWorkloadSpec:
    name: str
    preset: str
    n_reads: int
    seed: int
    n_contigs: int            # contigs of <= 1 Mb (the fork's index limit)
    contig_len: int
    read_len: int
    err: float = 0.10
    p_repeat: float = 0.03
    max_occ: int = 50         # opt->mid_occ
    flag: int = 0
    b: int = 14               # mi->b


SEQ_CONFIGS = {
    # BASELINE.json configs[1] as sequences: 100k x 10 kb reads at 10 % error against 100 x 1 Mb of synthetic reference
    "map-ont-100k-seq": SeqWorkloadSpec("map-ont-100k-seq", "map-ont", 100_000, _SEED + 21, 100, 1_000_000, 10_000),
    "map-ont-small-seq": SeqWorkloadSpec("map-ont-small-seq", "map-ont", 2_000, _SEED + 22, 6, 300_000, 10_000),
}


class SeqWorkload:
    """The genome (bytes), a sampler of reads, and -- given the genome's minimizers -- its index streams."""

    def __init__(self, spec: SeqWorkloadSpec):
        L = lib()
        if not hasattr(L, "_seq_bound"):
            L.wlg_genome.argtypes = [C.POINTER(_GenomeSpec), C.c_void_p]
            L.wlg_genome.restype = None
            L.wlg_reads_count.restype = C.c_int64
            L.wlg_reads_count.argtypes = [C.POINTER(_GenomeSpec), C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]
            L.wlg_reads_fill.restype = None
            L.wlg_reads_fill.argtypes = [C.POINTER(_GenomeSpec), C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]
            L.wls_build_sites.restype = C.c_void_p
            L.wls_build_sites.argtypes = [C.c_int32, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p]
            L.wls_free.argtypes = [C.c_void_p]
            L.wls_free.restype = None
            L.wls_streams.argtypes = [C.c_void_p] + [C.POINTER(C.c_void_p), C.POINTER(C.c_int64)] * 4 + [C.POINTER(C.c_int64)] * 2
            L.wls_streams.restype = None
            L._seq_bound = True
        self.spec = spec
        self._cs = _GenomeSpec(spec.seed, spec.n_contigs, spec.contig_len, spec.p_repeat, spec.read_len, spec.err)
        self.genome = np.zeros(spec.n_contigs * spec.contig_len, np.uint8)
        L.wlg_genome(C.byref(self._cs), self.genome.ctypes.data)
        self._idx = None
        self.streams = []

    def contig_offsets(self) -> np.ndarray:
        return np.arange(self.spec.n_contigs + 1, dtype=np.int64) * self.spec.contig_len

    def build_index(self, mv_off: np.ndarray, mv: np.ndarray, k: int):
        """Index streams from the genome's minimizers (mm_sketch of contig c with rid = c, as mm_idx_gen does, index.c:340-416):
        key = x >> 8, position packed refid 21 | lastpos 21 | strand 1 | rank 21 with rank = the contig's own number."""
        x, y = mv[:, 0], mv[:, 1]
        rid = (y >> np.uint64(32)) & np.uint64(0x1fffff)
        pos = (rid << np.uint64(43)) | (((y & np.uint64(0xffffffff)) >> np.uint64(1)) & np.uint64(0x1fffff)) << np.uint64(22) | \
              (y & np.uint64(1)) << np.uint64(21) | rid
        key = np.ascontiguousarray(x >> np.uint64(8))
        pos = np.ascontiguousarray(pos)
        L = lib()
        self._idx = L.wls_build_sites(self.spec.b, 2 * k, key.shape[0], key.ctypes.data, pos.ctypes.data)
        if not self._idx:
            raise MemoryError("wls_build_sites failed")
        ptr = [C.c_void_p() for _ in range(4)]
        size = [C.c_int64() for _ in range(4)]
        nk, mo = C.c_int64(), C.c_int64()
        args = []
        for p, s in zip(ptr, size):
            args += [C.byref(p), C.byref(s)]
        L.wls_streams(self._idx, *args, C.byref(nk), C.byref(mo))
        self.streams = []
        for p, s in zip(ptr, size):
            n = int(s.value)
            self.streams.append(np.frombuffer((C.c_char * max(n, 1)).from_address(p.value), np.uint8, n) if n else np.zeros(0, np.uint8))
        self.n_keys, self.max_occ_seen = int(nk.value), int(mo.value)
        return self.streams

    def reads_count(self, first: int, n_reads: int):
        off = np.zeros(n_reads + 1, np.int64)
        bid = np.zeros(max(n_reads, 1), np.uint32)
        lib().wlg_reads_count(C.byref(self._cs), self.genome.ctypes.data, first, n_reads, off.ctypes.data, bid.ctypes.data)
        return off, bid[:n_reads]

    def reads_fill(self, first: int, off: np.ndarray, out: np.ndarray) -> None:
        assert out.dtype == np.uint8 and out.size >= int(off[-1])
        lib().wlg_reads_fill(C.byref(self._cs), self.genome.ctypes.data, first, off.shape[0] - 1, off.ctypes.data, out.ctypes.data)

    def close(self):
        if self._idx:
            self.streams = []
            lib().wls_free(self._idx)
            self._idx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

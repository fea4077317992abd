"""Shared helpers for the parity tests: adversarial anchor generators and result comparison."""
from __future__ import annotations

import numpy as np


def pack(strand, rid, rpos, span, qpos, seg=0, flags=0):
    x = (np.asarray(strand, np.uint64) << np.uint64(63)) | (np.asarray(rid, np.uint64) << np.uint64(32)) | np.asarray(rpos, np.uint64)
    y = (np.asarray(seg, np.uint64) << np.uint64(48)) | (np.asarray(flags, np.uint64) << np.uint64(40)) \
        | (np.asarray(span, np.uint64) << np.uint64(32)) | (np.asarray(qpos, np.uint64) & np.uint64(0xffffffff))
    return np.stack([x, y], axis=1).astype(np.uint64)


def sort_by_x(a, rng=None):
    """Sort by x only; equal-x anchors keep a random (but fixed) relative order, like the reference's unstable sort."""
    if rng is not None:
        a = a[rng.permutation(a.shape[0])]
    return a[np.argsort(a[:, 0], kind="stable")]


def random_dense(rng, n, rmax=3000, qmax=3000, n_rid=2, span_lo=1, span_hi=40, segs=1):
    """Small coordinate space: many ties, dr==0, dq<=0, overlapping windows, equal scores."""
    a = pack(rng.integers(0, 2, n), rng.integers(0, n_rid, n), rng.integers(0, rmax, n),
             rng.integers(span_lo, span_hi + 1, n), rng.integers(0, qmax, n), seg=rng.integers(0, segs, n))
    return sort_by_x(a, rng)


def random_colinear(rng, n, noise=0.2, step=8, jitter=3, span=15, band=300):
    """One noisy diagonal plus off-diagonal repeats (deep windows, exercises max_skip)."""
    q = np.cumsum(rng.integers(1, step + 1, n))
    r = q + np.cumsum(rng.integers(-jitter, jitter + 1, n)) + 10000
    off = rng.random(n) < noise
    r = np.where(off, r + rng.integers(-band, band + 1, n), r)
    a = pack(np.zeros(n, np.uint64), np.zeros(n, np.uint64), np.maximum(r, 0), np.full(n, span), q)
    return sort_by_x(a, rng)


def interleaved_diagonals(rng, n, k, step=3, gap=700, span=15, drop=0.1):
    """k parallel diagonals more than bw apart, visited round-robin along x: for an anchor on one diagonal the
    predecessors on the other k-1 are SKIPped (dd > bw), so every 32-predecessor chunk holds only a few scored lanes,
    the marks t[p[j]] land k anchors back (beyond the chunk for k > 32) and the max_skip break comes ~26*k deep."""
    r = 20000 + np.cumsum(rng.integers(1, step + 1, n))
    d = np.arange(n) % k
    q = r - 20000 + d * gap + rng.integers(-2, 3, n)
    keep = rng.random(n) >= drop
    a = pack(np.zeros(n, np.uint64), np.zeros(n, np.uint64), r, np.full(n, span), q)[keep]
    return sort_by_x(a, rng)


def many_chain_reads(rng, n_reads=6, n_chains=150):
    """Reads with > 64 short chains, many of which start at the same x: the final x-order hits ksort's unstable path."""
    reads = []
    for _ in range(n_reads):
        parts = []
        for c in range(n_chains):
            n = int(rng.integers(3, 9))
            q0 = int(rng.integers(0, 1000))
            r0 = int(rng.integers(0, 40)) * 100000
            q = q0 + np.arange(n) * 20
            parts.append(np.stack([np.full(n, (c % 7) << 32, np.uint64) + np.uint64(r0) + (np.arange(n) * 20).astype(np.uint64),
                                   (np.uint64(15) << np.uint64(32)) | q.astype(np.uint64)], axis=1))
        a = np.concatenate(parts).astype(np.uint64)
        a = a[rng.permutation(a.shape[0])]
        reads.append(a[np.argsort(a[:, 0], kind="stable")])
    return reads


def concat_reads(reads):
    offs = np.zeros(len(reads) + 1, np.int64)
    for i, r in enumerate(reads):
        offs[i + 1] = offs[i] + r.shape[0]
    a = np.concatenate([r.reshape(-1, 2) for r in reads], axis=0) if reads else np.zeros((0, 2), np.uint64)
    return offs, np.ascontiguousarray(a, dtype=np.uint64)


def assert_same(got, want, what=""):
    for name in ("n_u", "u_off", "b_off", "u", "b"):
        g, w = getattr(got, name), getattr(want, name)
        if g.shape != w.shape or not np.array_equal(g, w):
            bad = ""
            if name == "n_u" and g.shape == w.shape:
                idx = np.nonzero(g != w)[0]
                bad = f" first differing reads {idx[:8].tolist()} got {g[idx[:8]].tolist()} want {w[idx[:8]].tolist()}"
            raise AssertionError(f"{what}: field {name} differs (got shape {g.shape}, want {w.shape}).{bad}")

"""Shared inputs for the post-processing tests (row f3): chains of workload samples and fuzzed chain sets."""
import numpy as np


def fuzz_chains(rng, n_chains, qlen=20000, n_rid=2, joinable=0.4, max_cnt=30):
    """A read's chains as mm_chain_dp returns them: u[] (score<<32 | cnt) and b[] (each chain ascending, chains in the x
    order of their first anchors).  Query intervals overlap a lot (parents / secondaries), and a share of the chains continues
    the previous one colinearly on the same strand / reference within the join distances (mm_join_long candidates)."""
    chains = []
    prev = None
    for _ in range(n_chains):
        cnt = int(rng.integers(3, max_cnt))
        span = int(rng.choice([15, 15, 15, 19, 11]))
        if prev is not None and rng.random() < joinable:
            strand, rid, x0, q0 = prev
            gap_x = int(rng.integers(1, 4000 if rng.random() < 0.7 else 24000))
            gap_q = gap_x + int(rng.integers(-1500, 1500)) if rng.random() < 0.8 else int(rng.integers(1, 24000))
            x, q = x0 + max(gap_x, 1), q0 + max(gap_q, 1)
        else:
            strand, rid = int(rng.integers(2)), int(rng.integers(n_rid))
            x, q = int(rng.integers(100, 200000)), int(rng.integers(span, max(qlen - 6000, span + 1)))
        xs, qs = [x], [q]
        for _ in range(cnt - 1):
            step = int(rng.integers(1, 400))
            xs.append(xs[-1] + step + int(rng.integers(0, 3)))
            qs.append(qs[-1] + step + int(rng.integers(0, 3)))
        prev = (strand, rid, xs[-1], qs[-1])
        score = int(rng.integers(40, 400)) if rng.random() < 0.5 else int(rng.integers(400, 6000))
        if rng.random() < 0.1 and chains:
            score = chains[-1][0]                                  # equal scores: the hash decides the order
        a = np.zeros((cnt, 2), np.uint64)
        a[:, 0] = (np.uint64(strand) << np.uint64(63)) | (np.uint64(rid) << np.uint64(32)) | np.array(xs, np.uint64)
        a[:, 1] = (np.uint64(span) << np.uint64(32)) | np.array(qs, np.uint64)
        chains.append((score, a))
    chains.sort(key=lambda c: int(c[1][0, 0]))                     # chain.c:199-212: by the x of the first anchor
    u = np.array([(c[0] << 32) | c[1].shape[0] for c in chains], np.uint64)
    b = np.concatenate([c[1] for c in chains]) if chains else np.zeros((0, 2), np.uint64)
    qmax = int((b[:, 1] & np.uint64(0xffffffff)).max()) + 50 if chains else qlen
    return u, b, max(qlen, qmax)


def read_hash(r, seed=11):
    """stand-in for the read-name hash of map.c:347-349 (any 32-bit value; it only breaks score ties)"""
    x = (r * 2654435761 + seed * 40503) & 0xffffffff
    x ^= x >> 15
    return (x * 2246822519) & 0xffffffff

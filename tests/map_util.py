"""A small synthetic genome, its index streams (built from the ORACLE's sketch of the contigs) and reads sampled from it, for
the composed pipeline (bases in, regions out)."""
import numpy as np


def make_genome(rng, n_contigs=5, lo=20000, hi=40000, repeat=True):
    acgt = np.frombuffer(b"ACGT", np.uint8)
    contigs = [acgt[rng.integers(0, 4, int(rng.integers(lo, hi)))].copy() for _ in range(n_contigs)]
    if repeat:                                             # a 3 kb segment present in three contigs (one reverse-complemented): multi-hit minimizers, secondaries
        seg = contigs[0][2000:5000].copy()
        contigs[1][7000:10000] = seg
        comp = np.zeros(256, np.uint8); comp[[65, 67, 71, 84]] = [84, 71, 67, 65]
        contigs[2][1000:4000] = comp[seg[::-1]]
    return [c.tobytes() for c in contigs]


def index_streams(contigs, w, k, hpc, b=8):
    """{minimizer: [packed positions]} from the oracle's mm_sketch of every contig -> the fork's four index streams"""
    from oracle import post as opost, seed as oseed
    table = {}
    for rid, seq in enumerate(contigs):
        mv = opost.sketch(seq, w, k, rid=rid, is_hpc=bool(hpc))
        for x, y in mv:
            x, y = int(x), int(y)
            table.setdefault(x >> 8, []).append(oseed.pack_position(rid, (y & 0xffffffff) >> 1, y & 1, rid))
    return oseed.build_streams(table, b=b, rng=np.random.default_rng(3), load=0.6, deleted=0.05), table


def sample_reads(rng, contigs, n_reads, lo=800, hi=9000, err=0.08):
    comp = np.zeros(256, np.uint8); comp[[65, 67, 71, 84]] = [84, 71, 67, 65]
    comp[ord("N")] = ord("N")
    acgt = np.frombuffer(b"ACGT", np.uint8)
    reads = []
    for t in range(n_reads):
        c = np.frombuffer(contigs[int(rng.integers(len(contigs)))], np.uint8)
        L = int(rng.integers(lo, min(hi, c.size - 1)))
        p = int(rng.integers(0, c.size - L))
        s = c[p:p + L].copy()
        if t % 11 == 5:                                    # chimeric: two loci joined
            c2 = np.frombuffer(contigs[int(rng.integers(len(contigs)))], np.uint8)
            p2 = int(rng.integers(0, c2.size - L // 2))
            s = np.concatenate([s[:L // 2], c2[p2:p2 + L // 2]])
        keep = rng.random(s.size) >= err / 3               # deletions
        s = s[keep]
        sub = rng.random(s.size) < err / 3                 # substitutions
        s[sub] = acgt[rng.integers(0, 4, int(sub.sum()))]
        ins = np.nonzero(rng.random(s.size) < err / 3)[0]  # insertions
        s = np.insert(s, ins, acgt[rng.integers(0, 4, ins.size)])
        if t % 7 == 3:
            q = int(rng.integers(0, s.size)); s[q:q + int(rng.integers(1, 30))] = ord("N")
        if rng.random() < 0.5:
            s = comp[s[::-1]]
        if t % 13 == 7:
            s = acgt[rng.integers(0, 4, int(rng.integers(50, 3000)))].copy()     # unrelated read: no chain
        reads.append(s.tobytes())
    reads.append(b"")                                      # an empty read
    reads.append(b"ACGTACGTAC")                            # shorter than w + k
    return reads

#!/usr/bin/env python
"""sim_reads.py <ref.fa> <reads.fa> [--ref-len N] [--contigs C] [--reads R] [--read-len L] [--err E] [--seed S]
G-seq generator of SURVEY.md 8(d): a random reference (uniform ACGT) cut into contigs, reads sampled uniformly with a
random strand and a per-base error rate split 1/3 substitutions, 1/3 insertions, 1/3 deletions."""
import argparse

import numpy as np

COMP = {"A": "T", "C": "G", "G": "C", "T": "A"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("ref"); ap.add_argument("reads_out")
    ap.add_argument("--ref-len", type=int, default=300000); ap.add_argument("--contigs", type=int, default=3)
    ap.add_argument("--n-reads", type=int, default=120); ap.add_argument("--read-len", type=int, default=5000)
    ap.add_argument("--err", type=float, default=0.10); ap.add_argument("--seed", type=int, default=20261104)
    a = ap.parse_args()
    rng = np.random.default_rng(a.seed)
    contigs = ["".join(rng.choice(list("ACGT"), a.ref_len // a.contigs)) for _ in range(a.contigs)]
    # a diverged repeat inside contig 0 so that some windows hold off-diagonal predecessors
    c0 = list(contigs[0]); seg = c0[1000:4000]
    for k in range(len(seg)):
        if rng.random() < 0.03:
            seg[k] = "ACGT"[int(rng.integers(4))]
    c0[20000:23000] = seg; contigs[0] = "".join(c0)
    with open(a.ref, "w") as fh:
        for k, c in enumerate(contigs):
            fh.write(f">ctg{k}\n{c}\n")
    with open(a.reads_out, "w") as fh:
        for r in range(a.n_reads):
            c = contigs[int(rng.integers(len(contigs)))]
            st = int(rng.integers(0, max(len(c) - a.read_len, 1)))
            s = c[st:st + a.read_len]
            if rng.random() < 0.5:
                s = "".join(COMP[x] for x in reversed(s))
            out = []
            for ch in s:
                u = rng.random()
                if u < a.err / 3:
                    out.append("ACGT"[int(rng.integers(4))])
                elif u < 2 * a.err / 3:
                    out.append(ch); out.append("ACGT"[int(rng.integers(4))])
                elif u < a.err:
                    continue
                else:
                    out.append(ch)
            fh.write(f">read{r}\n{''.join(out)}\n")


if __name__ == "__main__":
    main()

"""Test sequences for the sketching checks (row f4): every branch of mm_sketch (sketch.c:77-143)."""
import numpy as np

SETTINGS = [(10, 15, 0), (5, 15, 0), (10, 19, 1), (5, 19, 1), (19, 19, 0), (11, 21, 0), (3, 6, 0), (1, 4, 0), (4, 8, 1), (50, 28, 1),
            (255, 11, 0), (16, 14, 0), (7, 2, 1),
            (3, 7, 0), (1, 5, 0), (19, 13, 0), (10, 1, 0), (200, 15, 0), (2, 3, 0)]   # (w, k, is_hpc); the presets use the first six;
                                                        # no HPC with an odd k <= 15 runs the specialised kernel (k_sketch_odd)


def sequences(rng, n=24, max_len=6000):
    """random, ambiguous, homopolymer-rich, tandem-repeat (equal k-mers inside one window), palindromic (k-mer == its reverse
    complement: even k only), lower-case / non-ACGT bytes, very short and empty-window cases"""
    out = []
    acgt = np.frombuffer(b"ACGT", np.uint8)
    for t in range(n):
        L = int(rng.integers(1, max_len)) if t % 6 else int(rng.integers(1, 60))
        s = acgt[rng.integers(0, 4, L)].copy()
        kind = t % 8
        if kind == 1:                                       # stretches of N and single Ns
            for _ in range(int(rng.integers(1, 12))):
                p = int(rng.integers(0, L)); s[p:p + int(rng.integers(1, 40))] = ord("N")
        elif kind == 2:                                     # homopolymer runs, some longer than 255
            for _ in range(int(rng.integers(1, 20))):
                p = int(rng.integers(0, L)); s[p:p + int(rng.integers(2, 30 if rng.random() < 0.8 else 400))] = acgt[int(rng.integers(4))]
        elif kind == 3:                                     # tandem repeats: identical k-mers at several window positions
            unit = acgt[rng.integers(0, 4, int(rng.integers(1, 25)))]
            p = int(rng.integers(0, max(L - 1, 1))); m = min(L - p, int(rng.integers(20, 600)))
            s[p:p + m] = np.resize(unit, m)
        elif kind == 4:                                     # reverse-complement palindromes
            half = acgt[rng.integers(0, 4, int(rng.integers(2, 20)))]
            comp = {65: 84, 67: 71, 71: 67, 84: 65}
            pal = np.concatenate([half, np.array([comp[int(b)] for b in half[::-1]], np.uint8)])
            for _ in range(int(rng.integers(1, 8))):
                p = int(rng.integers(0, L)); m = min(L - p, pal.size); s[p:p + m] = pal[:m]
        elif kind == 5:                                     # lower case, U, other bytes, raw 0..3 codes
            sel = rng.random(L) < 0.3
            s[sel] = np.frombuffer(b"acgtuURYKMn-*", np.uint8)[rng.integers(0, 13, int(sel.sum()))]
            s[rng.random(L) < 0.02] = int(rng.integers(0, 4))
        elif kind == 6:                                     # two-letter low complexity
            s = np.frombuffer(b"AT", np.uint8)[rng.integers(0, 2, L)].copy()
        out.append(s.tobytes())
    return out

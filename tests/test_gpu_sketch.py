"""Minimizer sketching on the device (include/mm2sketch_b200.h) against the oracle (oracle/sketch_oracle.c,
pinned to the reference's sketch.c in tests/test_sketch_oracle.py), through the C ABI: the same minimizers in the same order."""
import os

import numpy as np
import pytest

from sketch_util import SETTINGS, sequences

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    from minimap2_chaindp_b200 import api
    c = api.ChainContext(device=0, profile=True)
    yield c
    c.close()


def _check(ctx, seqs, w, k, hpc, rid=None):
    from oracle import post as opost
    from minimap2_chaindp_b200 import sketch
    mv_off, mv = sketch.sketch_batch(ctx, seqs, w, k, bool(hpc), rid)
    total = 0
    for t, seq in enumerate(seqs):
        want = opost.sketch(seq, w, k, rid=0 if rid is None else int(rid[t]), is_hpc=bool(hpc)) if len(seq) else np.zeros((0, 2), np.uint64)
        got = mv[mv_off[t]:mv_off[t + 1]]
        if not np.array_equal(got, want):
            first = next((i for i in range(min(len(got), len(want))) if not np.array_equal(got[i], want[i])), min(len(got), len(want)))
            raise AssertionError(f"w={w} k={k} hpc={hpc} read {t} (len {len(seq)}): {len(got)} vs {len(want)} minimizers, first difference at {first}")
        total += want.shape[0]
    return total


@pytest.mark.parametrize("w,k,hpc", SETTINGS)
def test_every_branch_of_mm_sketch(ctx, w, k, hpc):
    rng = np.random.default_rng(1000 * w + 10 * k + hpc)
    seqs = sequences(rng, n=32, max_len=5000) + [b"", b"A", b"N" * 700, b"ACGT" * 600, b"A" * 3000 + b"C" + b"G" * 1500]
    assert _check(ctx, seqs, w, k, hpc, rid=np.arange(len(seqs)) % 7) > 0


@pytest.mark.parametrize("w,k,hpc", SETTINGS[:6] + [(200, 15, 0), (3, 7, 0), (255, 11, 0)])
def test_reads_across_many_tiles(ctx, w, k, hpc):
    """10-60 kb reads (20-120 tiles of 512 bases): the carried ring, k-mer context and counters"""
    rng = np.random.default_rng(77 + w + k + hpc)
    acgt = np.frombuffer(b"ACGT", np.uint8)
    seqs = []
    for t in range(12):
        L = int(rng.integers(10000, 60000))
        s = acgt[rng.integers(0, 4, L)].copy()
        if t % 3 == 1:
            for _ in range(30):
                p = int(rng.integers(0, L)); s[p:p + int(rng.integers(1, 20))] = ord("N")
        if t % 3 == 2:
            for _ in range(200):
                p = int(rng.integers(0, L)); s[p:p + int(rng.integers(2, 40))] = acgt[int(rng.integers(4))]
            p = int(rng.integers(0, L - 2000)); s[p:p + 1500] = np.resize(acgt[rng.integers(0, 4, 7)], 1500)
        seqs.append(s.tobytes())
    # tile boundaries inside homopolymer runs, N stretches and repeats
    seqs.append(b"ACGTTGCA" * 64 + b"A" * 600 + b"N" * 513 + b"CA" * 2000)
    assert _check(ctx, seqs, w, k, hpc) > 1000


def test_fixture_sequences(ctx):
    d = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "test")
    seqs = []
    for fn in ("MT-human.fa", "MT-orang.fa", "t-inv.fa", "q-inv.fa"):
        path = os.path.join(d, fn)
        if os.path.exists(path):
            seqs.append(b"".join(l.strip() for l in open(path, "rb") if not l.startswith(b">")))
    if not seqs:
        pytest.skip("fixture sequences not present (copied at build time from /root/reference/test)")
    for w, k, hpc in SETTINGS[:6]:
        assert _check(ctx, seqs, w, k, hpc) > 100


def test_errors_and_empty_batch(ctx):
    from minimap2_chaindp_b200 import api, sketch
    mv_off, mv = sketch.sketch_batch(ctx, [], 10, 15)
    assert mv_off.tolist() == [0] and mv.shape[0] == 0
    for w, k in ((0, 15), (256, 15), (10, 0), (10, 29)):
        with pytest.raises(api.ChainError) as ei:
            sketch.sketch_batch(ctx, [b"ACGT"], w, k)
        assert ei.value.code == api.ERR_ARG
    assert sketch.sketch_stage_ms(ctx)[1] >= 0

"""Pins the sketching restatement (oracle/sketch_oracle.c) to the reference: mm_sketch of the compiled,
unmodified sketch.c (oracle/_ref/libmm2ref_post.so) must return the same minimizers in the same order."""
import numpy as np
import pytest

from oracle import post
from sketch_util import SETTINGS, sequences

needs_ref = pytest.mark.skipif(not post.have_ref(), reason="oracle/_ref/libmm2ref_post.so not built (needs /root/reference at build time)")


@needs_ref
@pytest.mark.parametrize("w,k,hpc", SETTINGS)
def test_restatement_equals_mm_sketch(w, k, hpc):
    rng = np.random.default_rng(1000 * w + 10 * k + hpc)
    total = 0
    for t, seq in enumerate(sequences(rng)):
        got = post.sketch(seq, w, k, rid=t, is_hpc=bool(hpc))
        want = post.ref_sketch(seq, w, k, rid=t, is_hpc=bool(hpc))
        assert np.array_equal(got, want), f"w={w} k={k} hpc={hpc} sequence {t} (len {len(seq)})"
        assert got.shape[0] <= len(seq)
        total += got.shape[0]
    assert total > 0


@needs_ref
def test_fixture_sequences():
    import os
    d = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "test")
    n = 0
    for fn in ("MT-human.fa", "MT-orang.fa", "t-inv.fa", "q-inv.fa", "t2.fa", "q2.fa"):
        path = os.path.join(d, fn)
        if not os.path.exists(path):
            continue
        seq = b"".join(l.strip() for l in open(path, "rb") if not l.startswith(b">"))
        for w, k, hpc in SETTINGS[:6]:
            assert np.array_equal(post.sketch(seq, w, k, 0, bool(hpc)), post.ref_sketch(seq, w, k, 0, bool(hpc))), (fn, w, k, hpc)
            n += 1
    assert n > 0

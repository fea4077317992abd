"""Pins the oracle (our CPU restatement) against the reference's own chain.c compiled unmodified (oracle/_ref)."""
import numpy as np
import pytest

import oracle
from minimap2_chaindp_b200 import workload
from util import concat_reads, many_chain_reads, random_colinear, random_dense

needs_ref = pytest.mark.skipif(not oracle.have_ref(), reason="oracle/_ref not built (no /root/reference on this box)")


def _check_batch(prm, offs, a, what):
    got = oracle.chain_batch(prm, offs, a)
    ref = oracle.ref_chain_batch(prm, offs, a, n_threads=4)
    assert got.same_as(ref), what
    return got


@needs_ref
@pytest.mark.parametrize("name,n_reads", [("map-ont-100k", 1500), ("ava-ont-50k", 60), ("map-pb-ul-20k", 6)])
def test_workload_samples(name, n_reads):
    spec = workload.CONFIGS[name]
    offs, a = workload.generate(spec, 0, n_reads)
    res = _check_batch(oracle.params(spec.preset), offs, a, name)
    assert res.u_off[-1] > 0 and res.visited > 0


@needs_ref
@pytest.mark.parametrize("seed", range(4))
def test_adversarial_dense(seed):
    rng = np.random.default_rng(1000 + seed)
    reads = [random_dense(rng, int(n)) for n in rng.integers(0, 400, 150)]
    offs, a = concat_reads(reads)
    for over in (dict(), dict(min_cnt=1, min_sc=0), dict(max_skip=0), dict(max_skip=3, bw=50), dict(max_dist_x=200, max_dist_y=150)):
        _check_batch(oracle.params("map-ont", **over), offs, a, f"dense {seed} {over}")


@needs_ref
@pytest.mark.parametrize("seed", range(3))
def test_deep_windows(seed):
    rng = np.random.default_rng(2000 + seed)
    reads = [random_colinear(rng, int(n), noise=float(rng.random()) * 0.6, step=int(rng.integers(1, 12)))
             for n in rng.integers(2, 3000, 30)]
    offs, a = concat_reads(reads)
    for over in (dict(), dict(max_skip=5), dict(max_skip=100), dict(n_segs=5000), dict(bw=2000)):
        _check_batch(oracle.params("map-ont", **over), offs, a, f"deep {seed} {over}")


@needs_ref
def test_generic_path():
    rng = np.random.default_rng(3000)
    reads = [random_dense(rng, int(n), segs=2) for n in rng.integers(0, 300, 100)]
    offs, a = concat_reads(reads)
    for preset, over in (("sr", dict()), ("splice", dict()), ("map-ont", dict(n_segs=2)), ("splice", dict(bw=300))):
        _check_batch(oracle.params(preset, **over), offs, a, f"generic {preset} {over}")


@needs_ref
def test_split_form_matches_monolithic():
    """mm_chain_dp == mm_chain_dp_fpga + mm_chain_dp_bottom, and our compaction == new_seed[]."""
    spec = workload.CONFIGS["map-ont-100k"]
    offs, a = workload.generate(spec, 300, 40)
    prm = oracle.params(spec.preset)
    for r in range(40):
        ar = a[offs[r]:offs[r + 1]]
        u0, b0 = oracle.ref_chain_dp(prm, ar)
        rec = oracle.ref_chain_dp_fpga(prm, ar)
        u1, b1 = oracle.ref_chain_dp_bottom(prm, rec)
        assert np.array_equal(u0, u1) and np.array_equal(b0, b1)
        o = oracle.fill(prm, ar)
        orec, newid = oracle.compact(prm.min_sc, ar, o.f, o.p, o.v)
        assert np.array_equal(orec, rec)
        u2, b2 = oracle.bottom(prm, orec)
        assert np.array_equal(u0, u2) and np.array_equal(b0, b2)
        # new_seed.f / .p expose f[] and p[] of every kept anchor: pins the instrumented fill
        keep = newid >= 0
        assert np.array_equal(rec["f"][newid[keep]], o.f[keep])


@needs_ref
@pytest.mark.parametrize("n", [0, 1, 2, 63, 64, 65, 200, 5000])
def test_radix_sorts(n):
    rng = np.random.default_rng(n)
    for bits in (3, 9, 17, 40, 64):
        z = rng.integers(0, 2 ** min(bits, 63), n, dtype=np.uint64) << np.uint64(max(0, bits - 63))
        assert np.array_equal(oracle.radix_sort_64(z), oracle.ref_radix_sort_64(z))
        w = np.stack([z, np.arange(n, dtype=np.uint64)], axis=1)
        assert np.array_equal(oracle.radix_sort_128x(w), oracle.ref_radix_sort_128x(w)), (n, bits)


@needs_ref
def test_many_chains_unstable_order():
    reads = many_chain_reads(np.random.default_rng(11))
    offs, a = concat_reads(reads)
    res = _check_batch(oracle.params("map-ont", min_sc=10, min_cnt=2), offs, a, "many chains")
    assert int(res.n_u.max()) > 64


def test_empty_read():
    prm = oracle.params("map-ont")
    u, b = oracle.chain_dp(prm, np.zeros((0, 2), np.uint64))
    assert u.shape[0] == 0 and b.shape[0] == 0
    if oracle.have_ref():
        u, b = oracle.ref_chain_dp(prm, np.zeros((0, 2), np.uint64))
        assert u.shape[0] == 0

"""Parity of the CUDA path against the oracle, through the C ABI (bit-exact: integer/index work)."""
import numpy as np
import pytest

import oracle
from minimap2_chaindp_b200 import api, workload
from util import assert_same, concat_reads, interleaved_diagonals, many_chain_reads, random_colinear, random_dense

pytestmark = pytest.mark.gpu


def _both(ctx, preset, offs, a, **over):
    got = ctx.chain_batch(api.params(preset, **over), offs, a)
    want = oracle.chain_batch(oracle.params(preset, **over), offs, a)
    return got, want


def test_fill_arrays_match_oracle(ctx):
    """f[], p[], v[] and the new_seed records (the fork's device/host boundary, chain.c:218-327)."""
    spec = workload.CONFIGS["map-ont-100k"]
    offs, a = workload.generate(spec, 0, 300)
    prm = api.params(spec.preset)
    rec_off, rec, f, p, v = ctx.fill_batch(prm, offs, a, want_fpv=True)
    oprm = oracle.params(spec.preset)
    for r in range(offs.shape[0] - 1):
        ar = a[offs[r]:offs[r + 1]]
        o = oracle.fill(oprm, ar)
        s = slice(offs[r], offs[r + 1])
        assert np.array_equal(f[s], o.f), f"read {r}: f differs"
        assert np.array_equal(p[s], o.p), f"read {r}: p differs"
        assert np.array_equal(v[s], o.v), f"read {r}: v differs"
        orec, _ = oracle.compact(oprm.min_sc, ar, o.f, o.p, o.v)
        got = rec[rec_off[r]:rec_off[r + 1]]
        assert got.shape == orec.shape and np.array_equal(got, orec), f"read {r}: new_seed records differ"


@pytest.mark.parametrize("name,n_reads", [("map-ont-100k", 3000), ("ava-ont-50k", 150), ("map-pb-ul-20k", 12)])
def test_configs_sample(ctx, name, n_reads):
    spec = workload.CONFIGS[name]
    offs, a = workload.generate(spec, 0, n_reads)
    got, want = _both(ctx, spec.preset, offs, a)
    assert_same(got, want, name)
    assert got.n_units > 0


@pytest.mark.parametrize("seed", range(6))
def test_adversarial_dense(ctx, seed):
    """Tiny coordinate space: ties everywhere, dr == 0, dq <= 0, many equal scores, many short chains."""
    rng = np.random.default_rng(1000 + seed)
    reads = [random_dense(rng, int(n)) for n in rng.integers(0, 400, 200)]
    offs, a = concat_reads(reads)
    for over in (dict(), dict(min_cnt=1, min_sc=0), dict(max_skip=0), dict(max_skip=3, bw=50), dict(max_dist_x=200, max_dist_y=150)):
        got, want = _both(ctx, "map-ont", offs, a, **over)
        assert_same(got, want, f"dense seed {seed} {over}")


@pytest.mark.parametrize("seed", range(4))
def test_deep_windows(ctx, seed):
    """Noisy diagonals with repeats: windows of hundreds of predecessors, the skip/break logic across chunks."""
    rng = np.random.default_rng(2000 + seed)
    reads = [random_colinear(rng, int(n), noise=float(rng.random()) * 0.6, step=int(rng.integers(1, 12)))
             for n in rng.integers(2, 3000, 40)]
    offs, a = concat_reads(reads)
    for over in (dict(), dict(max_skip=5), dict(max_skip=100), dict(n_segs=5000), dict(max_iter=50), dict(bw=2000)):
        got, want = _both(ctx, "map-ont", offs, a, **over)
        assert_same(got, want, f"deep seed {seed} {over}")


@pytest.mark.parametrize("ring", [64, -1, 128, 256])
def test_ring_depths(ring, monkeypatch):
    """The fill kernel's ring depth (MM2B200_RING; -1 = the 64-anchor ring with the batched loads of deep windows, what
    sub-batches of long reads get; 128 / 256 use them too): chunks 1..RING/32-1 come from
    shared memory with their marks in the ring's tag words, deeper ones from global memory with marks in t[] -- four chunks
    at a time in the batched variants, with the marks reloaded when an earlier chunk of the batch let a mark leave.  Dense and
    sparse windows, break budgets that cross several chunks, and the visited-pair counter, all against the oracle."""
    monkeypatch.setenv("MM2B200_RING", str(ring))
    rng = np.random.default_rng(7700 + abs(ring))
    reads = [random_colinear(rng, int(n), noise=float(rng.random()) * 0.6, step=int(rng.integers(1, 6))) for n in rng.integers(2, 2500, 20)]
    reads += [interleaved_diagonals(rng, int(n), k, step=int(rng.integers(1, 4))) for k in (3, 9, 40) for n in rng.integers(200, 3000, 4)]
    offs, a = concat_reads(reads)
    with api.ChainContext(device=0) as c, api.ChainContext(device=0, count_visited=True) as cc:
        for over in (dict(), dict(max_skip=3), dict(max_skip=200), dict(bw=100), dict(max_dist_x=20000, max_dist_y=20000),
                     dict(max_iter=300), dict(min_cnt=1, min_sc=0)):
            want = oracle.chain_batch(oracle.params("map-ont", **over), offs, a)
            assert_same(c.chain_batch(api.params("map-ont", **over), offs, a), want, f"ring {ring} {over}")
            got = cc.chain_batch(api.params("map-ont", **over), offs, a)
            assert_same(got, want, f"ring {ring} counted {over}")
            assert got.n_visited == want.visited, (ring, over)


def test_long_reads_pick_the_deep_ring(ctx):
    """A sub-batch of long reads (>= 4096 anchors per read on average) runs the long-read variant of the fill (batched deep loads)."""
    spec = workload.CONFIGS["map-pb-ul-20k"]
    offs, a = workload.generate(spec, 5, 3)
    got = ctx.chain_batch(api.params(spec.preset), offs, a)
    assert_same(got, oracle.chain_batch(oracle.params(spec.preset), offs, a), "map-pb reads, deep ring")


@pytest.mark.parametrize("k", [2, 5, 9, 20, 33, 50])
def test_sparse_chunks(ctx, k):
    """Chunks with 0-4 scored lanes (scalar path of the fill kernel), marks and breaks that cross chunk boundaries."""
    rng = np.random.default_rng(4000 + k)
    reads = [interleaved_diagonals(rng, int(n), k, step=int(rng.integers(1, 4))) for n in rng.integers(2, 4000, 24)]
    offs, a = concat_reads(reads)
    with api.ChainContext(device=0, count_visited=True) as c:
        for over in (dict(), dict(max_skip=3), dict(max_skip=60), dict(bw=100), dict(max_dist_x=20000, max_dist_y=20000), dict(max_iter=300)):
            want = oracle.chain_batch(oracle.params("map-ont", **over), offs, a)
            got = ctx.chain_batch(api.params("map-ont", **over), offs, a)
            assert_same(got, want, f"sparse k={k} {over}")
            got = c.chain_batch(api.params("map-ont", **over), offs, a)
            assert_same(got, want, f"sparse counted k={k} {over}")
            assert got.n_visited == want.visited, (k, over)


def test_generic_path_cdna_and_segments(ctx):
    """is_cdna and multi-segment anchors take the generic kernel (chain.c:71-77)."""
    rng = np.random.default_rng(3000)
    reads = [random_dense(rng, int(n), segs=2) for n in rng.integers(0, 300, 100)]
    offs, a = concat_reads(reads)
    for preset, over in (("sr", dict()), ("splice", dict()), ("map-ont", dict(n_segs=2)), ("splice", dict(bw=300))):
        got, want = _both(ctx, preset, offs, a, **over)
        assert_same(got, want, f"generic {preset} {over}")
    reads = [random_colinear(rng, 500, noise=0.3) for _ in range(20)]
    offs, a = concat_reads(reads)
    got, want = _both(ctx, "splice", offs, a)
    assert_same(got, want, "splice colinear")


def test_visited_pair_counter():
    """The device-side count of reference-visited (i,j) pairs (the metric's numerator) equals the oracle's."""
    with api.ChainContext(device=0, count_visited=True) as c:
        for name, n_reads in (("map-ont-100k", 800), ("ava-ont-50k", 40), ("map-pb-ul-20k", 4)):
            spec = workload.CONFIGS[name]
            offs, a = workload.generate(spec, 11, n_reads)
            got = c.chain_batch(api.params(spec.preset), offs, a)
            want = oracle.chain_batch(oracle.params(spec.preset), offs, a)
            assert_same(got, want, name)
            assert got.n_visited == want.visited, name
        rng = np.random.default_rng(5)
        reads = [random_colinear(rng, int(n), noise=0.4, step=3) for n in rng.integers(2, 2000, 30)]
        offs, a = concat_reads(reads)
        for over in (dict(), dict(max_skip=2), dict(bw=2000), dict(max_iter=40)):
            got = c.chain_batch(api.params("map-ont", **over), offs, a)
            want = oracle.chain_batch(oracle.params("map-ont", **over), offs, a)
            assert_same(got, want, f"count {over}")
            assert got.n_visited == want.visited, over


def test_empty_and_tiny(ctx):
    rng = np.random.default_rng(7)
    reads = [np.zeros((0, 2), np.uint64), random_dense(rng, 1), np.zeros((0, 2), np.uint64), random_dense(rng, 2),
             random_colinear(rng, 3), np.zeros((0, 2), np.uint64)]
    offs, a = concat_reads(reads)
    got, want = _both(ctx, "map-ont", offs, a, min_cnt=1, min_sc=0)
    assert_same(got, want, "tiny")
    got = ctx.chain_batch(api.params("map-ont"), np.zeros(1, np.int64), np.zeros((0, 2), np.uint64))
    assert got.n_u.shape[0] == 0 and got.u.shape[0] == 0


def test_many_chains_unstable_order(ctx):
    """> 64 kept chains with equal first-anchor x: the x-order must replay ksort's unstable radix sort."""
    reads = many_chain_reads(np.random.default_rng(11))
    offs, a = concat_reads(reads)
    got, want = _both(ctx, "map-ont", offs, a, min_sc=10, min_cnt=2)
    assert int(want.n_u.max()) > 64
    assert_same(got, want, "many chains")


def test_sub_batching_is_invisible():
    """Small slots force many sub-batches and slot reuse; results must not depend on the split."""
    spec = workload.CONFIGS["map-ont-100k"]
    offs, a = workload.generate(spec, 100, 1200)
    want = oracle.chain_batch(oracle.params(spec.preset), offs, a)
    for slots, cap in ((1, 20000), (2, 50000), (3, 7000)):
        with api.ChainContext(device=0, n_slots=slots, slot_anchors=cap) as c:
            got = c.chain_batch(api.params(spec.preset), offs, a)
            assert got.n_sub_batches > 3
            assert_same(got, want, f"slots={slots} cap={cap}")
            got = c.chain_batch(api.params(spec.preset), offs, a)      # arena reuse
            assert_same(got, want, f"slots={slots} cap={cap} (second call)")


def test_pinned_input_path(ctx):
    spec = workload.CONFIGS["map-ont-100k"]
    offs = workload.count(spec, 5000, 500)
    pa = api.PinnedArray((int(offs[-1]), 2), np.uint64)
    workload.fill(spec, offs, pa.array, 5000)
    got = ctx.chain_batch(api.params(spec.preset), offs, pa.array)
    want = oracle.chain_batch(oracle.params(spec.preset), offs, pa.array.copy())
    assert_same(got, want, "pinned")
    pa.free()


def test_reference_entry_points(ctx):
    """mm_chain_dp / mm_chain_dp_fpga / mm_chain_dp_bottom symbols with the reference's calling convention."""
    spec = workload.CONFIGS["map-ont-100k"]
    offs, a = workload.generate(spec, 42, 8)
    prm, oprm = api.params(spec.preset), oracle.params(spec.preset)
    for r in range(8):
        ar = a[offs[r]:offs[r + 1]]
        wu, wb = oracle.chain_dp(oprm, ar)
        u, b = api.mm_chain_dp(prm, ar)
        assert np.array_equal(u, wu) and np.array_equal(b, wb)
        rec = api.mm_chain_dp_fpga(prm, ar)
        o = oracle.fill(oprm, ar)
        orec, _ = oracle.compact(oprm.min_sc, ar, o.f, o.p, o.v)
        assert np.array_equal(rec, orec)
        u2, b2 = api.mm_chain_dp_bottom(prm, rec)
        assert np.array_equal(u2, wu) and np.array_equal(b2, wb)
    u, b = api.mm_chain_dp(prm, np.zeros((0, 2), np.uint64))        # n == 0 (test/t2.fa, q2.fa)
    assert u.shape[0] == 0 and b.shape[0] == 0


def test_concurrent_callers_are_batched(ctx):
    """Many host threads calling mm_chain_dp at once (kt_for workers): every caller gets its own read's answer."""
    import threading
    spec = workload.CONFIGS["map-ont-100k"]
    offs, a = workload.generate(spec, 7000, 64)
    prm, oprm = api.params(spec.preset), oracle.params(spec.preset)
    out = [None] * 64

    def work(r):
        out[r] = api.mm_chain_dp(prm, a[offs[r]:offs[r + 1]])

    th = [threading.Thread(target=work, args=(r,)) for r in range(64)]
    [t.start() for t in th]
    [t.join() for t in th]
    for r in range(64):
        wu, wb = oracle.chain_dp(oprm, a[offs[r]:offs[r + 1]])
        assert np.array_equal(out[r][0], wu) and np.array_equal(out[r][1], wb), f"read {r}"


def test_concurrent_split_entry_points_are_batched(ctx):
    """The fork's LIVE call sites (map.c:525,941 mm_chain_dp_fpga; map.c:837 mm_chain_dp_bottom) from 128 host threads:
    every caller gets its own read's records / chains, and the dispatcher combined the calls into fewer device batches."""
    import threading
    spec = workload.CONFIGS["map-ont-100k"]
    n = 128
    offs, a = workload.generate(spec, 9000, n)
    prm, oprm = api.params(spec.preset), oracle.params(spec.preset)
    out = [None] * n
    b0, c0 = api.entry_stats()

    def work(r):
        rec = api.mm_chain_dp_fpga(prm, a[offs[r]:offs[r + 1]])
        out[r] = (rec, api.mm_chain_dp_bottom(prm, rec))

    th = [threading.Thread(target=work, args=(r,)) for r in range(n)]
    [t.start() for t in th]
    [t.join() for t in th]
    b1, c1 = api.entry_stats()
    assert c1 - c0 == 2 * n
    assert b1 - b0 < 2 * n, "no two calls were ever combined"
    for r in range(n):
        ar = a[offs[r]:offs[r + 1]]
        o = oracle.fill(oprm, ar)
        orec, _ = oracle.compact(oprm.min_sc, ar, o.f, o.p, o.v)
        wu, wb = oracle.chain_dp(oprm, ar)
        assert np.array_equal(out[r][0], orec), f"records of read {r}"
        assert np.array_equal(out[r][1][0], wu) and np.array_equal(out[r][1][1], wb), f"chains of read {r}"


def test_mixed_entry_points_hammer(ctx):
    """mm_chain_dp from "worker" threads while mm_chain_dp_bottom runs from "result" threads (the layout of map.c:837/858
    with the rechain block enabled) and a third group uses other parameters: the dispatcher serialises the batches on its
    private context and nobody reads another call's arena."""
    import threading
    spec = workload.CONFIGS["map-ont-100k"]
    n = 48
    offs, a = workload.generate(spec, 12000, n)
    prm, oprm = api.params(spec.preset), oracle.params(spec.preset)
    prm2, oprm2 = api.params(spec.preset, min_cnt=1, min_sc=0), oracle.params(spec.preset, min_cnt=1, min_sc=0)
    recs = []
    for r in range(n):
        ar = a[offs[r]:offs[r + 1]]
        o = oracle.fill(oprm, ar)
        recs.append(oracle.compact(oprm.min_sc, ar, o.f, o.p, o.v)[0])
    want = [oracle.chain_dp(oprm, a[offs[r]:offs[r + 1]]) for r in range(n)]
    want2 = [oracle.chain_dp(oprm2, a[offs[r]:offs[r + 1]]) for r in range(n)]
    bad = []

    def chain(r):
        for _ in range(3):
            u, b = api.mm_chain_dp(prm, a[offs[r]:offs[r + 1]])
            if not (np.array_equal(u, want[r][0]) and np.array_equal(b, want[r][1])):
                bad.append(("mm_chain_dp", r))

    def bottom(r):
        for _ in range(3):
            u, b = api.mm_chain_dp_bottom(prm, recs[r])
            if not (np.array_equal(u, want[r][0]) and np.array_equal(b, want[r][1])):
                bad.append(("mm_chain_dp_bottom", r))

    def chain2(r):
        u, b = api.mm_chain_dp(prm2, a[offs[r]:offs[r + 1]])
        if not (np.array_equal(u, want2[r][0]) and np.array_equal(b, want2[r][1])):
            bad.append(("mm_chain_dp other params", r))

    th = [threading.Thread(target=f, args=(r,)) for r in range(n) for f in (chain, bottom, chain2)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not bad, bad[:5]


def test_max_skip_int_max(ctx):
    """max_skip = INT_MAX never breaks (chain.c:84); the budget arithmetic must not overflow."""
    rng = np.random.default_rng(5)
    reads = [random_colinear(rng, int(n), noise=0.3, step=5) for n in rng.integers(2, 400, 12)]
    offs, a = concat_reads(reads)
    got, want = _both(ctx, "map-ont", offs, a, max_skip=(1 << 31) - 1)
    assert_same(got, want, "max_skip INT_MAX")


def test_chain_device_then_other_stream(ctx):
    """mm2b200_chain_device returns before its work is done; a following batch call (other stream, may regrow the slot)
    must wait for it instead of freeing the workspace under the running kernels."""
    import torch
    spec = workload.CONFIGS["map-ont-100k"]
    offs, a = workload.generate(spec, 300, 3000)
    prm, oprm = api.params(spec.preset), oracle.params(spec.preset)
    dev = torch.device("cuda", 0)
    d_a = torch.from_numpy(a.view(np.int64)).to(dev)
    d_offs = torch.from_numpy(offs).to(dev)
    side = torch.cuda.Stream()
    with api.ChainContext(device=0, n_slots=1, slot_anchors=1 << 16) as c2:
        for _ in range(3):
            out = c2.chain_device(prm, 3000, int(offs[-1]), d_offs.data_ptr(), d_a.data_ptr(), side.cuda_stream)
            got = c2.chain_batch(prm, offs[:201], a[:offs[200]])      # different stream, smaller sub-batches
            want = oracle.chain_batch(oprm, offs[:201], a[:offs[200]])
            assert_same(got, want, "batch after chain_device")
        side.synchronize()
        got = c2.chain_batch(prm, offs, a)
        assert_same(got, oracle.chain_batch(oprm, offs, a), "full batch after chain_device")


def test_argument_errors(ctx):
    """Bad arguments come back as error codes (never abort, never compute)."""
    a = np.zeros((10, 2), np.uint64)
    with pytest.raises(api.ChainError) as ei:
        ctx.chain_batch(api.params("map-ont"), np.array([0, 7, 5, 10], np.int64), a)        # offs not non-decreasing
    assert ei.value.code == api.ERR_ARG
    with pytest.raises(api.ChainError) as ei:
        ctx.chain_batch(api.params("map-ont", max_dist_x=-1), np.array([0, 10], np.int64), a)
    assert ei.value.code == api.ERR_ARG
    with pytest.raises(api.ChainError) as ei:
        ctx.chain_batch(api.params("map-ont", max_iter=-5), np.array([0, 10], np.int64), a)
    assert ei.value.code == api.ERR_ARG
    with pytest.raises(api.ChainError) as ei:
        api.ChainContext(device=99)
    assert ei.value.code == api.ERR_CUDA
    # the context is still usable after an error
    got = ctx.chain_batch(api.params("map-ont"), np.array([0, 10], np.int64), a)
    assert got.n_u.tolist() == [0]


def test_wide_window_variant(ctx):
    """max_dist_x >= 2^26 switches to the prefix-of-lanes window test (WIDE)."""
    rng = np.random.default_rng(21)
    reads = [random_colinear(rng, int(n), noise=0.3, step=5) for n in rng.integers(2, 800, 25)]
    offs, a = concat_reads(reads)
    for over in (dict(max_dist_x=1 << 26, max_dist_y=1 << 26), dict(max_dist_x=1 << 27, max_dist_y=1 << 27), dict(max_dist_x=(1 << 31) - 1, max_dist_y=5000, bw=700)):
        got, want = _both(ctx, "map-ont", offs, a, **over)
        assert_same(got, want, f"wide {over}")

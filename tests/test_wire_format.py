"""The host batcher's 8-byte upload format (include/mm2chain_b200.h: mm2b200_set_upload / mm2b200_wire_encode).

CPU part: the encoder alone against a numpy restatement of the device decoder (lossless round trip, refusal of anchors
that do not fit).  GPU part: chains are identical with the format on and off, with several encoder threads and
sub-batches, and when a sub-batch has to fall back to the raw upload."""
import numpy as np
import pytest

from minimap2_chaindp_b200 import api, workload
from util import assert_same, concat_reads, pack, random_colinear, random_dense, sort_by_x


def _roundtrip(a):
    a = np.asarray(a, dtype=np.uint64).reshape(-1, 2)
    enc = api.wire_encode(a)
    assert enc is not None
    packed, tab, d = enc
    nblk = (len(a) + api.WIRE_BLOCK - 1) // api.WIRE_BLOCK
    assert len(packed) == len(a) and len(tab) == nblk and len(d) <= 2 * len(a)
    assert np.array_equal(api.wire_decode(packed, tab, d), a)
    return enc


@pytest.mark.parametrize("seed", range(4))
def test_roundtrip_random(seed):
    rng = np.random.default_rng(seed)
    _roundtrip(sort_by_x(random_dense(rng, int(rng.integers(1, 9000)), n_rid=int(rng.integers(1, 9)))))
    _roundtrip(random_colinear(rng, int(rng.integers(2, 5000))))


def test_roundtrip_workloads():
    for name, n_reads in (("map-ont-100k", 40), ("ava-ont-50k", 4), ("map-pb-ul-20k", 2)):
        offs, a = workload.generate(workload.CONFIGS[name], 0, n_reads)
        packed, tab, d = _roundtrip(a)
        assert len(d) * 4 < 2 * len(packed)        # dictionaries stay below 2 B/anchor on the BASELINE workloads


def test_block_boundaries_and_empty():
    rng = np.random.default_rng(7)
    for n in (0, 1, api.WIRE_BLOCK - 1, api.WIRE_BLOCK, api.WIRE_BLOCK + 1, 3 * api.WIRE_BLOCK):
        _roundtrip(sort_by_x(random_dense(rng, n)) if n else np.zeros((0, 2), np.uint64))


def test_flags_segments_and_strands_survive():
    n = 3000
    k = np.arange(n, dtype=np.uint64)
    a = pack(k % 2, k % 5, 1000 + k * 7, 10 + k % 30, 5 + k * 3, seg=k % 3, flags=(k % 4) << 8)
    _roundtrip(a)


def test_refuses_wide_qpos():
    k = np.arange(2000, dtype=np.uint64)
    wide = pack(0 * k, 0 * k, 100 + k, 15 + 0 * k, (1 << 24) + k)
    assert api.wire_encode(wide) is None
    ok = pack(0 * k, 0 * k, 100 + k, 15 + 0 * k, (1 << 24) - 1 - k)
    _roundtrip(ok)
    # 32 distinct spans in a block take 5 index bits: qpos must then stay below 2^19
    assert api.wire_encode(pack(0 * k, 0 * k, 100 + k, 10 + k % 32, (1 << 19) + k)) is None
    _roundtrip(pack(0 * k, 0 * k, 100 + k, 10 + k % 32, (1 << 19) - 1 - k))
    _roundtrip(pack(0 * k, k, 100 + k, 15 + 0 * k, 5 + k))  # a new rid on every anchor: 256 x-entries per block


# ---------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_chains_identical_with_and_without_wire():
    import oracle
    spec = workload.CONFIGS["map-ont-100k"]
    offs, a = workload.generate(spec, 0, 1500)
    want = oracle.chain_batch(oracle.params(spec.preset), offs, a)
    # small sub-batches: several of them, each split over the encoder threads
    with api.ChainContext(device=0, slot_anchors=300_000, n_slots=3) as c:
        for wire, threads in ((True, 4), (True, 1), (False, 1), (True, 0)):
            c.set_upload(wire, threads)
            got = c.chain_batch(api.params(spec.preset), offs, a)
            assert_same(got, want, f"wire={wire} threads={threads}")
            nbytes, nwire = c.upload_stats()
            if wire:
                assert nwire == got.n_sub_batches and nbytes < 10 * len(a) + 16 * len(offs)
            else:
                assert nwire == 0 and nbytes >= 16 * len(a)


@pytest.mark.gpu
def test_wire_fallback_sub_batches(ctx):
    """Reads the format cannot carry (qpos >= 2^24) go up raw, next to reads that are encoded."""
    import oracle
    rng = np.random.default_rng(11)
    k = np.arange(1500, dtype=np.uint64)
    reads = [random_colinear(rng, 2500),
             pack(0 * k, 0 * k, 1000 + 20 * k, 15 + 0 * k, (1 << 24) + 20 * k),
             random_colinear(rng, 1800),
             sort_by_x(pack(k % 2, k % 700, 1000 + 20 * k, 10 + k % 32, (1 << 19) + 20 * k)),
             sort_by_x(random_dense(rng, 3000, n_rid=5))]
    offs, a = concat_reads(reads)
    want = oracle.chain_batch(oracle.params("map-ont"), offs, a)
    with api.ChainContext(device=0, slot_anchors=3000, n_slots=2) as c:     # one read per sub-batch
        c.set_upload(True, 2)
        got = c.chain_batch(api.params("map-ont"), offs, a)
        assert_same(got, want, "mixed wire/raw sub-batches")
        _, nwire = c.upload_stats()
        assert 0 < nwire < got.n_sub_batches
    assert_same(ctx.chain_batch(api.params("map-ont"), offs, a), want, "one sub-batch, raw")
    assert ctx.upload_stats()[1] == 0          # the format is opt-in


@pytest.mark.gpu
def test_fill_batch_through_wire():
    import oracle
    rng = np.random.default_rng(12)
    reads = [sort_by_x(random_dense(rng, int(n), n_rid=3)) for n in rng.integers(0, 3000, 30)]
    offs, a = concat_reads(reads)
    prm = api.params("map-ont")
    with api.ChainContext(device=0) as c:
        c.set_upload(True, 3)
        rec_off, rec, f, p, v = c.fill_batch(prm, offs, a, want_fpv=True)
        assert c.upload_stats()[1] >= 1
    oprm = oracle.params("map-ont")
    for r in range(len(reads)):
        o = oracle.fill(oprm, a[offs[r]:offs[r + 1]])
        s = slice(offs[r], offs[r + 1])
        assert np.array_equal(f[s], o.f) and np.array_equal(p[s], o.p) and np.array_equal(v[s], o.v)

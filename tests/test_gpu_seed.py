"""Seeding on the device against the oracle and the reference's own outputs.

Everything goes through the C ABI of include/mm2seed_b200.h.  Golden fixtures (tests/golden/seed_*.npz) hold what the
unmodified reference produced; synthetic index streams reach the cases they do not (long position lists, probe chains over
deleted slots, all-vs-all filters, x ties in reads of more than 64 anchors, ragged and empty reads).
"""
import glob
import os
import struct

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "seed_*.npz")))
IDS = [os.path.basename(p)[5:-4] for p in GOLD]


@pytest.fixture(scope="module")
def ctx():
    from minimap2_chaindp_b200 import api
    c = api.ChainContext(device=0, profile=True)
    yield c
    c.close()


def _golden(path, ctx):
    from minimap2_chaindp_b200 import api, seed
    z = np.load(path)
    ix = seed.SeedIndex(ctx)
    ix.load(z["idx_B"], seed.INDEX_B, chunk=100000)       # chunked like load_index_buf (main.c:179)
    ix.load(z["idx_H"], seed.INDEX_H, chunk=65536)
    ix.load(z["idx_V"], seed.INDEX_V)
    ix.load(z["idx_P"], seed.INDEX_P)
    ix.commit()
    bw, is_cdna, max_skip, min_sc, flag, max_occ = (int(v) for v in z["params"])
    meta = z["meta"]
    assert len(set(map(tuple, meta[:, :2]))) == 1          # one (gap_qry, gap_ref) per run
    gap_qry, gap_ref = int(meta[0, 0]), int(meta[0, 1])
    chain = api.params("map-ont", max_dist_x=gap_ref, max_dist_y=gap_qry, bw=bw, max_skip=max_skip, min_sc=min_sc,
                       is_cdna=is_cdna, n_segs=gap_qry)    # map.c:517
    prm = seed.seed_params(chain, flag, max_occ)
    bid = meta[:, 4] | (meta[:, 5] << 31)
    return z, ix, prm, bid, meta[:, 2]


@pytest.mark.parametrize("path", GOLD, ids=IDS)
def test_anchors_match_the_reference(path, ctx):
    z, ix, prm, bid, qlen = _golden(path, ctx)
    out = ix.seed_chain_batch(prm, z["mv_off"], z["mv"], bid, qlen, want_anchors=True)
    assert np.array_equal(out.a_off, z["a_off"])
    assert np.array_equal(out.a, z["a"])
    assert np.array_equal(out.rep_len, z["rep_len"].astype(np.int32))
    assert np.array_equal(out.mini_off, z["mini_off"]) and np.array_equal(out.mini_pos, z["mini"])
    assert ix.launch_count() >= 4
    ix.close()


@pytest.mark.parametrize("path", GOLD, ids=IDS)
def test_records_match_the_reference(path, ctx):
    z, ix, prm, bid, qlen = _golden(path, ctx)
    out = ix.seed_fill_batch(prm, z["mv_off"], z["mv"], bid, qlen)
    assert np.array_equal(out.rec_off, z["rec_off"])
    assert out.rec.tobytes() == z["rec"].tobytes()
    ix.close()


@pytest.mark.parametrize("path", GOLD, ids=IDS)
def test_chains_of_seeded_reads_match_the_oracle(path, ctx):
    import oracle
    z, ix, prm, bid, qlen = _golden(path, ctx)
    out = ix.seed_chain_batch(prm, z["mv_off"], z["mv"], bid, qlen)
    oprm = oracle.Params(**{k: getattr(prm.chain, k) for k, _ in oracle.Params._fields_})
    want = oracle.chain_batch(oprm, z["a_off"], z["a"])
    assert np.array_equal(out.chains.n_u, want.n_u) and np.array_equal(out.chains.u, want.u) and np.array_equal(out.chains.b, want.b)
    ix.close()


def _packet(z, r0, r1):
    """request packet of map.c:286-324 for reads [r0, r1)"""
    body = b""
    for r in range(r0, r1):
        gap_qry, gap_ref, qlen, read_id, rank, rflag, n_segs, b = (int(v) for v in z["meta"][r])
        mv = z["mv"][z["mv_off"][r]:z["mv_off"][r + 1]]
        task = struct.pack("<iiiiIIhbxQ28x", gap_qry, gap_ref, mv.shape[0], qlen, read_id, rank | rflag << 31, n_segs, b, 0)
        assert len(task) == 64
        raw = mv.tobytes()
        body += task + raw + b"\0" * (-len(raw) % 64)
    head = struct.pack("<IIHHBB50x", 0x1234abcd, 64 + len(body), 7, r1 - r0, 3, 1)
    assert len(head) == 64
    return head + body


@pytest.mark.parametrize("path", GOLD, ids=IDS)
def test_packet_level_response_matches_the_reference(path, ctx):
    z, ix, prm, bid, qlen = _golden(path, ctx)
    n = int(z["n_reads"][0])
    step = 16
    for r0 in range(0, n, step):
        r1 = min(n, r0 + step)
        resp = ix.work(prm, _packet(z, r0, r1))
        magic, size, tid, num, typ, lat = struct.unpack_from("<IIHHBB", resp, 0)
        assert (magic, tid, num, typ, lat) == (0x1234abcd, 7, r1 - r0, 3, 1) and size == len(resp)
        p = 64
        for r in range(r0, r1):
            err, read_id, sub, n_a, n_mini, rep_len = struct.unpack_from("<6I", resp, p)
            assert err == 0 and read_id == int(z["meta"][r, 3]) and rep_len == int(z["rep_len"][r])
            want_rec = z["rec"][24 * z["rec_off"][r]:24 * z["rec_off"][r + 1]].tobytes()
            assert n_a * 24 == len(want_rec) and resp[p + 64:p + 64 + len(want_rec)] == want_rec
            q = p + 64 + (24 * n_a + 63) // 64 * 64
            want_mini = z["mini"][z["mini_off"][r]:z["mini_off"][r + 1]].tobytes()
            assert n_mini * 8 == len(want_mini) and resp[q:q + len(want_mini)] == want_mini
            assert sub == 64 + (24 * n_a + 63) // 64 * 64 + (8 * n_mini + 63) // 64 * 64
            p += sub
        assert p == len(resp)
    ix.close()


# ---------------------------------------------------------------------------------------------------------
# synthetic index streams
# ---------------------------------------------------------------------------------------------------------
def _synthetic(rng, n_keys=4000, b=8, max_list=60, deleted=0.2, n_ref=6, n_rank=12):
    from oracle import seed as oseed
    table = {}
    while len(table) < n_keys:
        m = int(rng.integers(1 << 34))
        u = rng.random()
        n = 1 if u < 0.5 else int(rng.integers(2, 6)) if u < 0.9 else int(rng.integers(6, max_list))
        table[m] = [oseed.pack_position(int(rng.integers(n_ref)), int(rng.integers(1 << 16)), int(rng.integers(2)), int(rng.integers(n_rank)))
                    for _ in range(n)]
    return table, oseed.build_streams(table, b=b, rng=rng, load=0.7, deleted=deleted)


def _reads(rng, table, n_reads, lens, miss=0.2, repeat=0.15):
    """minimizer lists: keys of the table (some repeated at several query positions -> x ties, some adjacent -> tandem),
    unknown keys, ascending positions"""
    keys = np.array(list(table.keys()), dtype=np.uint64)
    mv, off, qlen = [], [0], []
    for r in range(n_reads):
        n = int(lens[r])
        ks = keys[rng.integers(0, keys.size, n)]
        if n > 4:
            hot = keys[rng.integers(0, keys.size, 3)]
            sel = rng.random(n) < repeat
            ks[sel] = hot[rng.integers(0, 3, int(sel.sum()))]
            dup = np.nonzero(rng.random(n - 1) < 0.05)[0]
            ks[dup + 1] = ks[dup]
        absent = rng.random(n) < miss
        ks[absent] = rng.integers(1 << 35, 1 << 36, int(absent.sum())).astype(np.uint64)
        pos = np.cumsum(rng.integers(1, 12, n)) + 20
        strand = rng.integers(0, 2, n)
        span = rng.integers(11, 20, n)
        m = np.zeros((n, 2), np.uint64)
        m[:, 0] = (ks << np.uint64(8)) | span.astype(np.uint64)
        m[:, 1] = (pos.astype(np.uint64) << np.uint64(1)) | strand.astype(np.uint64)
        mv.append(m); off.append(off[-1] + n); qlen.append(int(pos[-1]) + 40 if n else 50)
    return np.array(off, np.int64), (np.concatenate(mv) if mv else np.zeros((0, 2), np.uint64)), np.array(qlen, np.int32)


def _check_against_oracle(ctx, rng, oix, table, n_reads, lens, flag, max_occ, preset="map-ont", expect_exact=None):
    import oracle
    from oracle import seed as oseed
    from minimap2_chaindp_b200 import api, seed
    off, mv, qlen = _reads(rng, table, n_reads, lens)
    bid = (rng.integers(0, 12, n_reads).astype(np.int64) | (rng.integers(0, 2, n_reads).astype(np.int64) << 31))
    ix = seed.SeedIndex(ctx).load_streams(oix.B, oix.H, oix.V, oix.P)
    prm = seed.seed_params(api.params(preset), flag, max_occ)
    out = ix.seed_chain_batch(prm, off, mv, bid, qlen, want_anchors=True)
    want_a, want_off = [], [0]
    for r in range(n_reads):
        a, rep, mini = oseed.collect_seed_hits(oix, flag, max_occ, mv[off[r]:off[r + 1]], int(bid[r]), int(qlen[r]))
        assert np.array_equal(out.a[out.a_off[r]:out.a_off[r + 1]], a), f"anchors of read {r}"
        assert int(out.rep_len[r]) == rep, f"rep_len of read {r}"
        assert np.array_equal(out.mini_pos[out.mini_off[r]:out.mini_off[r + 1]], mini), f"mini_pos of read {r}"
        want_a.append(a); want_off.append(want_off[-1] + a.shape[0])
    alla = np.concatenate(want_a) if want_a else np.zeros((0, 2), np.uint64)
    want = oracle.chain_batch(oracle.params(preset), np.array(want_off, np.int64), alla)
    assert np.array_equal(out.chains.n_u, want.n_u) and np.array_equal(out.chains.u, want.u) and np.array_equal(out.chains.b, want.b)
    if expect_exact is not None:
        assert (out.n_exact_sorts > 0) == expect_exact
    ix.close()
    return out


@pytest.mark.parametrize("flag", [0, 0x1 | 0x2, 0x1, 0x100000, 0x200000, 0x2], ids=["plain", "ava", "no_diag", "for_only", "rev_only", "no_dual_alone"])
def test_synthetic_index_filters(ctx, flag):
    rng = np.random.default_rng(100 + (flag & 0xff) + (flag >> 20))
    table, oix = _synthetic(rng)
    lens = rng.integers(0, 400, 64)
    lens[5] = 0; lens[17] = 1; lens[63] = 0
    _check_against_oracle(ctx, rng, oix, table, 64, lens, flag, max_occ=20)


def test_ties_in_long_reads_follow_the_reference_sort(ctx):
    rng = np.random.default_rng(7)
    table, oix = _synthetic(rng, n_keys=600, b=5, max_list=12, deleted=0.3)
    lens = rng.integers(300, 2500, 24)
    out = _check_against_oracle(ctx, rng, oix, table, 24, lens, 0, max_occ=50, expect_exact=True)
    assert out.n_exact_sorts >= 10


def test_reads_longer_than_the_shared_memory_sort(ctx):
    rng = np.random.default_rng(8)
    table, oix = _synthetic(rng, n_keys=20000, b=10, max_list=10, deleted=0.0)
    lens = np.array([9000, 30, 12000, 0, 5000], np.int64)      # several thousand anchors: keys sorted in global memory
    out = _check_against_oracle(ctx, rng, oix, table, 5, lens, 0, max_occ=50)
    assert int(np.diff(out.a_off).max()) > 4096


def test_anchor_counts_around_the_sort_network_boundaries(ctx):
    """the anchor sort pads a read to a multiple of 32 keys and changes method at distances of 32: reads of every length from 0
    to 700 minimizers (and a few longer ones) give anchor counts on, just below and just above the multiples of 32"""
    rng = np.random.default_rng(11)
    table, oix = _synthetic(rng, n_keys=6000, b=8, max_list=8, deleted=0.1)
    lens = np.concatenate([np.arange(0, 700), rng.integers(700, 1800, 40)])
    out = _check_against_oracle(ctx, rng, oix, table, lens.size, lens, 0, max_occ=50)
    n_a = np.diff(out.a_off)
    assert {0, 1, 31}.issubset(set(int(x) % 32 for x in n_a)) and int(n_a.max()) > 2048 and ((n_a > 1024) & (n_a <= 2048)).any()


def test_repetitive_minimizers_and_rep_len(ctx):
    rng = np.random.default_rng(9)
    table, oix = _synthetic(rng, n_keys=1500, b=6, max_list=40)
    lens = rng.integers(50, 900, 40)
    for max_occ in (1, 2, 5, 1000):
        _check_against_oracle(ctx, rng, oix, table, 40, lens, 0, max_occ=max_occ)


def test_empty_batch_and_reads_without_hits(ctx):
    from minimap2_chaindp_b200 import api, seed
    rng = np.random.default_rng(10)
    table, oix = _synthetic(rng, n_keys=300, b=4)
    ix = seed.SeedIndex(ctx).load_streams(oix.B, oix.H, oix.V, oix.P)
    prm = seed.seed_params(api.params("map-ont"), 0, 10)
    out = ix.seed_chain_batch(prm, np.zeros(1, np.int64), np.zeros((0, 2), np.uint64), np.zeros(0, np.uint32), np.zeros(0, np.int32), True)
    assert out.n_anchors == 0 and out.chains.n_u.size == 0
    mv = np.zeros((50, 2), np.uint64)
    mv[:, 0] = (np.arange(50, dtype=np.uint64) + np.uint64(1 << 40)) << np.uint64(8) | np.uint64(15)
    mv[:, 1] = (np.arange(50, dtype=np.uint64) * np.uint64(5) + np.uint64(30)) << np.uint64(1)
    out = ix.seed_chain_batch(prm, np.array([0, 20, 20, 50], np.int64), mv, np.zeros(3, np.uint32), np.full(3, 400, np.int32), True)
    assert out.n_anchors == 0 and list(out.chains.n_u) == [0, 0, 0] and list(np.diff(out.mini_off)) == [20, 0, 30] and list(out.rep_len) == [0, 0, 0]
    ix.close()


def test_errors_are_reported(ctx):
    from minimap2_chaindp_b200 import api, seed
    ix = seed.SeedIndex(ctx)
    prm = seed.seed_params(api.params("map-ont"), 0, 10)
    with pytest.raises(api.ChainError) as ei:
        ix.seed_chain_batch(prm, np.zeros(1, np.int64), np.zeros((0, 2), np.uint64), [], [])
    assert ei.value.code == api.ERR_STATE                # not committed
    ix.load(np.zeros(48, np.uint8), seed.INDEX_B)
    with pytest.raises(api.ChainError) as ei:
        ix.commit()
    assert ei.value.code == api.ERR_ARG                  # 3 buckets: not a power of two
    with pytest.raises(api.ChainError):
        ix.load(b"x", 9)
    # a bucket that points outside H is caught on the device, not dereferenced
    ix2 = seed.SeedIndex(ctx)
    B = np.zeros((16, 2), np.uint64)
    B[:, 0] = np.uint64(8 << 24); B[:, 1] = np.uint64(1000 << 28)
    ix2.load_streams(B, np.zeros(64, np.uint8), np.zeros(64, np.uint8), np.zeros(0, np.uint8))
    mv = np.array([[(5 << 8) | 15, 40 << 1]], np.uint64)
    with pytest.raises(api.ChainError) as ei:
        ix2.seed_chain_batch(prm, np.array([0, 1], np.int64), mv, [0], [100])
    assert ei.value.code == api.ERR_ARG and "inconsistent" in str(ei.value)
    ix.close(); ix2.close()


def test_index_can_be_replaced_and_batches_repeat(ctx):
    rng = np.random.default_rng(12)
    from minimap2_chaindp_b200 import api, seed
    from oracle import seed as oseed
    ix = seed.SeedIndex(ctx)
    prm = seed.seed_params(api.params("map-ont"), 0, 30)
    for round_ in range(3):
        table, oix = _synthetic(rng, n_keys=500 + 700 * round_, b=4 + round_)
        ix.load_streams(oix.B, oix.H, oix.V, oix.P)
        assert ix.info()[0] == 4 + round_
        for _ in range(2):
            n = int(rng.integers(1, 30))
            off, mv, qlen = _reads(rng, table, n, rng.integers(0, 300, n))
            out = ix.seed_chain_batch(prm, off, mv, np.zeros(n, np.uint32), qlen, want_anchors=True)
            for r in range(n):
                a, rep, mini = oseed.collect_seed_hits(oix, 0, 30, mv[off[r]:off[r + 1]], 0, int(qlen[r]))
                assert np.array_equal(out.a[out.a_off[r]:out.a_off[r + 1]], a)
    ix.close()

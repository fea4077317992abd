"""The seeding oracle (oracle/seed_oracle.c) pinned against the unmodified reference.

tests/golden/seed_*.npz hold, per fixture run of the reference binary, everything its device receives (index streams,
parameters, request packets) and what its own collect_seed_hits + mm_chain_dp_fpga answered for every read.
"""
import glob
import os

import numpy as np
import pytest

import oracle
from oracle import seed as oseed

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "seed_*.npz")))


def load(path):
    z = np.load(path)
    ix = oseed.SeedIndex(z["idx_B"], z["idx_H"], z["idx_V"], z["idx_P"])
    return z, ix


def test_fixtures_present():
    assert len(GOLD) >= 4


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p)[5:-4] for p in GOLD])
def test_anchors_rep_len_mini_pos_match_reference(path):
    z, ix = load(path)
    bw, is_cdna, max_skip, min_sc, flag, max_occ = (int(v) for v in z["params"])
    n_anchor = 0
    for r in range(int(z["n_reads"][0])):
        gap_qry, gap_ref, qlen, read_id, rank, rflag, n_segs, b = (int(v) for v in z["meta"][r])
        assert b == ix.b
        mv = z["mv"][z["mv_off"][r]:z["mv_off"][r + 1]]
        bid = rank | rflag << 31
        a, rep_len, mini = oseed.collect_seed_hits(ix, flag, max_occ, mv, bid, qlen)
        want = z["a"][z["a_off"][r]:z["a_off"][r + 1]]
        assert a.shape == want.shape and np.array_equal(a, want), f"read {r}"
        assert rep_len == int(z["rep_len"][r])
        assert np.array_equal(mini, z["mini"][z["mini_off"][r]:z["mini_off"][r + 1]])
        n_anchor += a.shape[0]
    assert n_anchor == int(z["a_off"][-1])


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p)[5:-4] for p in GOLD])
def test_records_of_seeded_reads_match_reference(path):
    """seeding + fill + compaction = the response packet of fpga_work (map.c:484-575), n_segs = gap_qry included"""
    z, ix = load(path)
    bw, is_cdna, max_skip, min_sc, flag, max_occ = (int(v) for v in z["params"])
    for r in range(int(z["n_reads"][0])):
        gap_qry, gap_ref, qlen, read_id, rank, rflag, n_segs, b = (int(v) for v in z["meta"][r])
        mv = z["mv"][z["mv_off"][r]:z["mv_off"][r + 1]]
        a, _, _ = oseed.collect_seed_hits(ix, flag, max_occ, mv, rank | rflag << 31, qlen)
        prm = oracle.params("map-ont", max_dist_x=gap_ref, max_dist_y=gap_qry, bw=bw, max_skip=max_skip, min_sc=min_sc,
                            is_cdna=is_cdna, n_segs=gap_qry)      # map.c:517: n_segs is read from gap_qry
        fl = oracle.fill(prm, a)
        rec, _ = oracle.compact(min_sc, a, fl.f, fl.p, fl.v)
        want = z["rec"][24 * z["rec_off"][r]:24 * z["rec_off"][r + 1]].view(oracle.SEED_DTYPE)
        assert rec.shape == want.shape and rec.tobytes() == want.tobytes(), f"read {r}"


def test_lookup_agrees_with_index_contents():
    z, ix = load([p for p in GOLD if "gseq" in p][0])
    mv = z["mv"]
    hits = sum(len(ix.get(int(x) >> 8)) > 0 for x in mv[:4000, 0])
    assert hits > 100          # reads come from this reference: many minimizers are found
    assert ix.get((1 << 55) + 12345) == []


def test_synthetic_streams_round_trip():
    rng = np.random.default_rng(5)
    table = {}
    for _ in range(3000):
        m = int(rng.integers(1 << 30))
        n = 1 if rng.random() < 0.6 else int(rng.integers(2, 40))
        table[m] = [oseed.pack_position(int(rng.integers(8)), int(rng.integers(1 << 20)), int(rng.integers(2)), int(rng.integers(8)))
                    for _ in range(n)]
    for deleted in (0.0, 0.3):
        ix = oseed.build_streams(table, b=6, rng=rng, load=0.7, deleted=deleted)     # 64 buckets: long probe sequences
        for m, pos in list(table.items())[:1500]:
            assert ix.get(m) == pos
        for _ in range(500):
            m = int(rng.integers(1 << 30))
            if m not in table:
                assert ix.get(m) == []


def test_unstable_sort_of_tied_anchors_is_replayed():
    """reads of more than 64 anchors with x ties: radix_sort_128x is unstable there (ksort.h:116-151), and the oracle's
    sort is pinned to the reference's own (compiled) sort"""
    if not oracle.have_ref():
        pytest.skip("compiled reference absent")
    rng = np.random.default_rng(11)
    positions = [oseed.pack_position(1, int(p), 0, 0) for p in rng.integers(1, 1 << 20, 300)]
    table = {1000 + k: [positions[k]] for k in range(300)}
    ix = oseed.build_streams(table, b=4)
    # the same minimizers occur at several query positions -> equal x, different y
    ks = rng.integers(0, 300, 900)
    mv = np.zeros((900, 2), np.uint64)
    mv[:, 0] = ((1000 + ks).astype(np.uint64) << np.uint64(8)) | np.uint64(15)
    mv[:, 1] = (np.arange(900, dtype=np.uint64) * np.uint64(7) + np.uint64(20)) << np.uint64(1)
    unsorted, _, _ = oseed.collect_seed_hits(ix, 0, 10, mv, 0, 10000, sort=False)
    got, _, _ = oseed.collect_seed_hits(ix, 0, 10, mv, 0, 10000, sort=True)
    want = oracle.ref_radix_sort_128x(unsorted)
    assert np.array_equal(got, want)
    stable = unsorted[np.argsort(unsorted[:, 0], kind="stable")]
    assert not np.array_equal(stable, want)       # the case really is order-sensitive

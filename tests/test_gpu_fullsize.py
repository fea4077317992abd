"""BASELINE.json full-size runs, checked through size-independent properties (the oracle would take minutes here):
structure of the reference's output contract, invariance under sub-batching / sharding, and a random sample of reads
against the oracle."""
import hashlib

import numpy as np
import pytest

import oracle
from minimap2_chaindp_b200 import api, workload

pytestmark = pytest.mark.gpu


def _digest(res):
    h = hashlib.sha256()
    for name in ("n_u", "u_off", "b_off", "u", "b"):
        h.update(np.ascontiguousarray(getattr(res, name)).tobytes())
    return h.hexdigest()


def _check_contract(res, offs, a, prm):
    """Properties every mm_chain_dp result has (chain.c:163-214), whatever the input size."""
    n_reads = offs.shape[0] - 1
    assert res.n_u.shape[0] == n_reads and res.u_off[-1] == res.u.shape[0] and res.b_off[-1] == res.b.shape[0]
    assert np.array_equal(np.diff(res.u_off), res.n_u)
    cnt = (res.u & np.uint64(0xffffffff)).astype(np.int64)
    score = (res.u >> np.uint64(32)).astype(np.int64)
    assert cnt.min() >= prm.min_cnt and score.min() >= prm.min_sc          # chain.c:174-176 filters
    # per read: sum of chain lengths == anchors returned
    per_read = np.add.reduceat(cnt, res.u_off[:-1][res.n_u > 0]) if cnt.size else np.zeros(0, np.int64)
    assert np.array_equal(per_read, np.diff(res.b_off)[res.n_u > 0])
    # inside a chain anchors ascend in x and in query position (a chain follows predecessors with dr > 0, dq > 0)
    starts = np.concatenate([[0], np.cumsum(cnt)[:-1]])
    x, q = res.b[:, 0], (res.b[:, 1] & np.uint64(0xffffffff)).astype(np.int64)
    inner = np.ones(x.shape[0], bool)
    inner[starts] = False
    assert np.all(x[1:][inner[1:]] > x[:-1][inner[1:]])
    assert np.all(q[1:][inner[1:]] > q[:-1][inner[1:]])
    # chains of one read are ordered by the x of their first anchor (chain.c:199-212, needed by mm_join_long)
    first_x = x[starts]
    read_of_chain = np.repeat(np.arange(n_reads), res.n_u)
    same_read = read_of_chain[1:] == read_of_chain[:-1]
    assert np.all(first_x[1:][same_read] >= first_x[:-1][same_read])
    # every returned anchor is one of the read's input anchors
    rng = np.random.default_rng(0)
    for r in rng.choice(np.nonzero(res.n_u > 0)[0], 200, replace=False):
        inp = {(int(v[0]), int(v[1])) for v in a[offs[r]:offs[r + 1]]}
        out = res.b[res.b_off[r]:res.b_off[r + 1]]
        assert all((int(v[0]), int(v[1])) in inp for v in out)


def test_map_ont_100k_full_size():
    """BASELINE.json configs[1]: 100 k synthetic 10 kb reads, 77 M anchors, in one call."""
    spec = workload.CONFIGS["map-ont-100k"]
    offs, a = workload.generate(spec)
    prm = api.params(spec.preset)
    with api.ChainContext(device=0, count_visited=True) as c:
        res = c.chain_batch(prm, offs, a)
    assert res.n_sub_batches >= 8
    _check_contract(res, offs, a, prm)
    # invariance: other slot sizes / slot counts, and a 3-way shard merged in read order, give the same bytes
    want = _digest(res)
    with api.ChainContext(device=0, n_slots=2, slot_anchors=3_000_000) as c:
        assert _digest(c.chain_batch(prm, offs, a)) == want
    ctxs = [api.ChainContext(device=0, slot_anchors=5_000_000) for _ in range(3)]
    try:
        merged, _ = api.chain_batch_multi(ctxs, prm, offs, a)
        assert _digest(merged) == want
    finally:
        [c.close() for c in ctxs]
    # a random sample of reads against the oracle (and the visited-pair count on a contiguous slice)
    oprm = oracle.params(spec.preset)
    rng = np.random.default_rng(1)
    for r in rng.choice(offs.shape[0] - 1, 1500, replace=False):
        wu, wb = oracle.chain_dp(oprm, a[offs[r]:offs[r + 1]])
        u, b = res.read(int(r))
        assert np.array_equal(u, wu) and np.array_equal(b, wb), f"read {r}"
    assert res.n_visited > 0
    sl = slice(40000, 42001)
    o = offs[sl] - offs[sl][0]
    sample = oracle.chain_batch(oprm, o, a[offs[40000]:offs[42000]])
    with api.ChainContext(device=0, count_visited=True) as c:
        got = c.chain_batch(prm, o, a[offs[40000]:offs[42000]])
    assert got.n_visited == sample.visited


@pytest.mark.parametrize("name,n_reads,n_check", [("ava-ont-50k", 4000, 40), ("map-pb-ul-20k", 600, 6)])
def test_other_configs_large_slice(name, n_reads, n_check):
    """configs[2] and [3] at tens of millions of anchors: contract properties + sampled oracle parity."""
    spec = workload.CONFIGS[name]
    offs, a = workload.generate(spec, 0, n_reads)
    prm = api.params(spec.preset)
    with api.ChainContext(device=0) as c:
        res = c.chain_batch(prm, offs, a)
    _check_contract(res, offs, a, prm)
    oprm = oracle.params(spec.preset)
    rng = np.random.default_rng(2)
    for r in rng.choice(n_reads, n_check, replace=False):
        wu, wb = oracle.chain_dp(oprm, a[offs[r]:offs[r + 1]])
        u, b = res.read(int(r))
        assert np.array_equal(u, wu) and np.array_equal(b, wb), f"{name} read {r}"


def test_asm5_contig_scale():
    """configs[4]: one serial run of ~4 M anchors per contig (two contigs here); full oracle comparison."""
    spec = workload.CONFIGS["asm5-50mb"]
    offs, a = workload.generate(spec, 0, 2)
    assert int(np.diff(offs).min()) > 3_000_000
    prm = api.params(spec.preset)
    with api.ChainContext(device=0, count_visited=True) as c:
        res = c.chain_batch(prm, offs, a)
    want = oracle.chain_batch(oracle.params(spec.preset), offs, a)
    for name in ("n_u", "u_off", "b_off", "u", "b"):
        assert np.array_equal(getattr(res, name), getattr(want, name)), name
    assert res.n_visited == want.visited

"""Golden vectors from the reference binary itself (tests/golden/, regenerate with tests/golden/make_golden.sh):
the six fixture DP calls of SURVEY.md section 4, including BASELINE.json configs[0] (map-ont MT-human vs MT-orang)."""
import numpy as np
import pytest

import oracle
from golden_util import golden_calls, paf_pins

CALLS = golden_calls()


def test_fixture_inventory():
    n = sorted(c["a"].shape[0] for c in CALLS)
    assert n == [0, 0, 125, 223, 346, 682, 732]
    cfg0 = [c for c in CALLS if c["name"].startswith("map-ont_MT")][0]
    assert cfg0["params"] == dict(max_dist_x=5000, max_dist_y=5000, bw=500, max_skip=25, min_cnt=3, min_sc=40, is_cdna=0, n_segs=5000)
    assert [(int(x >> 32), int(x & 0xffffffff)) for x in cfg0["u"]] == [(3189, 342)]


def test_paf_lines_pin_chain_scores():
    """cm:i / s1:i of every PAF line the reference printed is one of the dumped chains (format.c:248)."""
    chains = {}
    for c in CALLS:
        chains.setdefault(c["name"].split("#")[0], set()).update((int(x >> 32), int(x & 0xffffffff)) for x in c["u"])
    pins = paf_pins()
    assert len(pins) == 4
    for name, s1, cm in pins:
        assert (s1, cm) in chains[name], (name, s1, cm)


@pytest.mark.parametrize("call", CALLS, ids=[c["name"] for c in CALLS])
def test_oracle_matches_golden(call):
    prm = oracle.params("map-ont", **call["params"])
    a = call["a"]
    o = oracle.fill(prm, a)
    rec, _ = oracle.compact(prm.min_sc, a, o.f, o.p, o.v)
    assert rec.shape == call["rec"].shape and np.array_equal(rec, call["rec"]), "new_seed records"
    u, b = oracle.bottom(prm, call["rec"])
    assert np.array_equal(u, call["u"]) and np.array_equal(b, call["b"]), "bottom"
    u, b = oracle.chain_dp(prm, a)
    assert np.array_equal(u, call["u"]) and np.array_equal(b, call["b"]), "chain_dp"


@pytest.mark.gpu
@pytest.mark.parametrize("call", CALLS, ids=[c["name"] for c in CALLS])
def test_gpu_matches_golden(ctx, call):
    from minimap2_chaindp_b200 import api
    prm = api.params("map-ont", **call["params"])
    a = call["a"]
    offs = np.array([0, a.shape[0]], np.int64)
    res = ctx.chain_batch(prm, offs, a)
    assert np.array_equal(res.u, call["u"]) and np.array_equal(res.b, call["b"]), "mm2b200_chain_batch"
    rec_off, rec = ctx.fill_batch(prm, offs, a)
    assert rec.shape == call["rec"].shape and np.array_equal(rec, call["rec"]), "mm2b200_fill_batch"
    res = ctx.bottom_batch(prm, np.array([0, call["rec"].shape[0]], np.int64), call["rec"])
    assert np.array_equal(res.u, call["u"]) and np.array_equal(res.b, call["b"]), "mm2b200_bottom_batch"
    # the reference's own entry points
    u, b = api.mm_chain_dp(prm, a)
    assert np.array_equal(u, call["u"]) and np.array_equal(b, call["b"]), "mm_chain_dp"
    if a.shape[0]:
        r2 = api.mm_chain_dp_fpga(prm, a)
        assert np.array_equal(r2, call["rec"]), "mm_chain_dp_fpga"
        u, b = api.mm_chain_dp_bottom(prm, call["rec"])
        assert np.array_equal(u, call["u"]) and np.array_equal(b, call["b"]), "mm_chain_dp_bottom"


@pytest.mark.gpu
def test_gpu_golden_as_one_batch(ctx):
    """All fixture calls that share a parameter set, submitted as one batch."""
    from minimap2_chaindp_b200 import api
    from util import concat_reads
    groups = {}
    for c in CALLS:
        groups.setdefault(tuple(sorted(c["params"].items())), []).append(c)
    for key, cs in groups.items():
        offs, a = concat_reads([c["a"] for c in cs])
        res = ctx.chain_batch(api.params("map-ont", **dict(key)), offs, a)
        for r, c in enumerate(cs):
            u, b = res.read(r)
            assert np.array_equal(u, c["u"]) and np.array_equal(b, c["b"]), c["name"]

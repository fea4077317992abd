"""Pins the post-processing restatement (oracle/post_oracle.c) to the reference: every field of every
mm_reg1_t and the rewritten anchors must equal what the reference's own hit.c (compiled unmodified into
oracle/_ref/libmm2ref_post.so) produces -- on the chains of the golden fixture runs, on workload samples and on fuzzed chain
sets that reach the parent / secondary / join / filter branches.  Also: the preset table used on the GPU box equals the
reference's mm_set_opt, and the PAF lines the reference printed for the fixture runs follow from the restatement."""
import os
import re

import numpy as np
import pytest

import oracle
from oracle import post
from minimap2_chaindp_b200 import workload
from golden_util import GOLDEN_DIR, golden_calls
from post_util import fuzz_chains, read_hash

needs_ref = pytest.mark.skipif(not post.have_ref(), reason="oracle/_ref/libmm2ref_post.so not built (needs /root/reference at build time)")


def _same(pp, h, qlen, u, b, tag):
    r1, a1, n1 = post.chain_post(pp, h, qlen, u, b)
    r2, a2, n2 = post.ref_chain_post(pp, h, qlen, u, b)
    assert r1.tobytes() == r2.tobytes(), f"{tag}: regs differ\n{r1}\n{r2}"
    assert n1 == n2 and np.array_equal(a1[:n1], a2[:n2]), f"{tag}: anchors differ"
    return r1


@needs_ref
def test_preset_table_equals_the_reference_options():
    for name, d in post.POST_PRESETS.items():
        p = post.ref_preset(name)
        for k, v in d.items():
            got = getattr(p, k)
            assert got == pytest.approx(v) if isinstance(v, float) else got == v, (name, k, got, v)
        assert p.min_cnt == oracle.PRESETS[name]["min_cnt"] and p.min_chain_score == oracle.PRESETS[name]["min_sc"]


@needs_ref
@pytest.mark.parametrize("name,preset,n_reads", [("map-ont-100k", "map-ont", 400), ("ava-ont-50k", "map-ont", 12), ("ava-ont-50k", "ava-ont", 12),
                                                 ("map-pb-ul-20k", "map-pb", 6)])
def test_workload_chains(name, preset, n_reads):
    spec = workload.CONFIGS[name]
    offs, a = workload.generate(spec, 0, n_reads)
    res = oracle.chain_batch(oracle.params(spec.preset), offs, a)
    pp = post.post_params(preset)
    n_regs = 0
    for r in range(n_reads):
        u = res.u[res.u_off[r]:res.u_off[r + 1]]
        b = res.b[res.b_off[r]:res.b_off[r + 1]]
        qlen = int((a[offs[r]:offs[r + 1], 1] & np.uint64(0xffffffff)).max()) + 100 if offs[r + 1] > offs[r] else 100
        n_regs += _same(pp, read_hash(r), qlen, u, b, f"{name} read {r}").shape[0]
    assert n_regs > 0


@needs_ref
@pytest.mark.parametrize("seed", range(6))
def test_fuzzed_chain_sets(seed):
    rng = np.random.default_rng(500 + seed)
    seen = dict(secondary=0, dropped=0, joined=0, unset=0)
    for t in range(150):
        n = int(rng.integers(1, 12)) if t % 5 else int(rng.integers(60, 160))
        u, b, qlen = fuzz_chains(rng, n, joinable=0.2 + 0.1 * (seed % 4))
        over = {}
        if seed == 1: over = dict(pri_ratio=0.3, best_n=2)
        if seed == 2: over = dict(mask_level=0.1, min_diff=200)
        if seed == 3: over = dict(flag=0x400)                       # no long join
        if seed == 4: over = dict(max_join_long=50000, max_join_short=20000, min_join_flank_sc=100)
        if seed == 5: over = dict(pri_ratio=0.0)
        pp = post.post_params("map-ont", **over)
        regs = _same(pp, read_hash(t, seed), qlen, u, b, f"seed {seed} case {t}")
        seen["secondary"] += int((regs["parent"] != regs["id"]).sum())
        seen["dropped"] += n - regs.shape[0]
        seen["joined"] += int((b[:, 1] != post.chain_post(pp, read_hash(t, seed), qlen, u, b)[1][:, 1]).sum())
        seen["unset"] += int((regs["parent"] < 0).sum())
    assert seen["secondary"] > 0 and seen["dropped"] > 0
    if seed in (0, 1, 2, 4):
        assert seen["joined"] > 0                                   # LONG_JOIN marks were written


@needs_ref
def test_degenerate_inputs():
    pp = post.post_params("map-ont")
    assert post.chain_post(pp, 1, 100, np.zeros(0, np.uint64), np.zeros((0, 2), np.uint64))[0].shape[0] == 0
    a = np.array([[100, (15 << 32) | 50], [130, (15 << 32) | 80], [160, (15 << 32) | 110]], np.uint64)
    for strand in (0, 1):
        b = a.copy(); b[:, 0] |= np.uint64(strand) << np.uint64(63)
        r = _same(pp, 7, 200, np.array([(45 << 32) | 3], np.uint64), b, f"single chain strand {strand}")
        assert r.shape[0] == 1 and r["parent"][0] == 0 and r["cnt"][0] == 3
    # two identical chains: the second is an identical hit and is dropped by mm_select_sub (hit.c:239)
    b2 = np.concatenate([a, a]); u2 = np.array([(45 << 32) | 3, (45 << 32) | 3], np.uint64)
    r = _same(pp, 9, 200, u2, b2, "identical chains")
    assert r.shape[0] == 1


def test_golden_paf_lines_follow_from_the_restatement():
    """PAF columns 3-4, 8-11 and the cm/s1/s2 tags of tests/golden/paf.txt (printed by the unmodified reference, no CIGAR)
    against the regions the restatement derives from the golden chains of the same runs."""
    lines, name = {}, None
    for line in open(os.path.join(GOLDEN_DIR, "paf.txt")):
        if line.startswith("## "):
            name = line[3:].split(":")[0]
            lines[name] = []
        elif line.strip():
            lines[name].append(line.rstrip("\n").split("\t"))
    checked = 0
    for call in golden_calls():
        run = call["name"].split("#")[0]
        if run.startswith("asm5") or call["u"].size == 0:           # asm5 fixture runs with -c: coordinates come from the base alignment
            continue
        preset = "ava-ont" if run.startswith("ava") else "map-ont"
        for f in lines[run]:
            qlen = int(f[1])
            regs, _, _ = post.chain_post(post.post_params(preset), 0, qlen, call["u"], call["b"])
            hit = [r for r in regs if int(r["qs"]) == int(f[2]) and int(r["qe"]) == int(f[3]) and int(r["rs"]) == int(f[7]) and int(r["re"]) == int(f[8])]
            if not hit:
                continue
            r = hit[0]
            tags = {t[:2]: t[5:] for t in f[12:]}
            assert int(r["mlen"]) == int(f[9]) and int(r["blen"]) == int(f[10])
            assert int(r["cnt"]) == int(tags["cm"]) and int(r["score"]) == int(tags["s1"])
            assert "s2" not in tags or int(r["subsc"]) == int(tags["s2"])
            assert "+-"[(int(r["bits"]) >> 10) & 1] == f[4]
            checked += 1
    assert checked >= 3

"""The rows of SURVEY.md 8(f) composed on the device (include/mm2map_b200.h): bases in, mm_reg1_t out -- against the same
composition of the oracles on the CPU: mm_sketch -> collect_seed_hits -> mm_chain_dp -> mm_gen_regs + chain_post.  Every field of
every region, rep_len and the stage counts (minimizers, anchors, chains) must agree."""
import numpy as np
import pytest

from map_util import index_streams, make_genome, sample_reads
from post_util import read_hash

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("preset,flag,max_occ,chunk,stages", [("map-ont", 0, 50, 0, 0), ("map-ont", 0, 50, 20000, 0), ("map-pb", 0, 50, 0, 0), ("map-pb", 0, 50, 1, 0),
                                                               ("ava-ont", 0x1 | 0x2 | 0x400 | 0x800000, 30, 50000, 0), ("asm5", 0, 10, 0, 0),
                                                               ("map-ont", 0, 50, 0, 2), ("map-pb", 0, 50, 0, 5), ("map-ont", 0, 50, 100000, 3), ("ava-ont", 0x1 | 0x2, 30, 0, 64)])
def test_bases_in_regions_out(preset, flag, max_occ, chunk, stages, monkeypatch):
    """chunk > 0: MM2B200_MAP_CHUNK bases per group (1 = every read its own group: the memory bound of the call); stages > 0: every
    group runs as a pipeline of that many chunks (MM2B200_MAP_STAGES, with MM2B200_MAP_STAGE_MIN lowered so that these small
    batches qualify): a chunk is seeded and chained while the next one is uploaded and sketched"""
    for name, v in (("MM2B200_MAP_CHUNK", chunk), ("MM2B200_MAP_STAGES", stages), ("MM2B200_MAP_STAGE_MIN", 1 if stages else 0)):
        if v:
            monkeypatch.setenv(name, str(v))
        else:
            monkeypatch.delenv(name, raising=False)
    if stages:
        monkeypatch.setenv("MM2B200_MAP_PART", "30000")                 # several upload parts per chunk
    import oracle
    from oracle import post as opost, seed as oseed
    from minimap2_chaindp_b200 import api, post, seed, sketch
    w, k, hpc = sketch.SKETCH_PRESETS[preset]
    rng = np.random.default_rng(sum(map(ord, preset)))
    contigs = make_genome(rng)
    oix, table = index_streams(contigs, w, k, hpc)
    reads = sample_reads(rng, contigs, 60)
    n = len(reads)
    bid = np.array([int(rng.integers(0, len(contigs))) | (int(rng.integers(0, 2)) << 31) for _ in range(n)], np.int64)
    hashes = np.array([read_hash(r, 5) for r in range(n)], np.uint32)
    chain_prm = api.params(preset)
    seq_off = np.concatenate([[0], np.cumsum([len(s) for s in reads])]).astype(np.int64)
    raw = np.frombuffer(b"".join(reads), np.uint8)
    with api.ChainContext(device=0, profile=True) as ctx:
        ix = seed.SeedIndex(ctx).load_streams(oix.B, oix.H, oix.V, oix.P)
        got = ix.map_regs_batch(sketch.SketchParams(w, k, hpc), seed.seed_params(chain_prm, flag, max_occ), post.post_params(preset),
                                seq_off, raw, bid, hashes)
        ix.close()
    # the same composition on the CPU
    oprm, opp = oracle.params(preset), opost.post_params(preset)
    n_mv = n_a = n_chains = n_regs = 0
    for r, s in enumerate(reads):
        mv = opost.sketch(s, w, k, rid=0, is_hpc=bool(hpc)) if len(s) else np.zeros((0, 2), np.uint64)
        a, rep, _ = oseed.collect_seed_hits(oix, flag, max_occ, mv, int(bid[r]), len(s))
        u, b = oracle.chain_dp(oprm, a)
        want, _, _ = opost.chain_post(opp, int(hashes[r]), len(s), u, b)
        g = got.regs[got.reg_off[r]:got.reg_off[r + 1]]
        assert int(got.rep_len[r]) == rep, f"{preset} read {r}: rep_len"
        assert g.tobytes() == want.tobytes(), f"{preset} read {r} ({len(s)} bases, {mv.shape[0]} minimizers, {a.shape[0]} anchors):\n{g}\n{want}"
        n_mv += mv.shape[0]; n_a += a.shape[0]; n_chains += u.shape[0]; n_regs += want.shape[0]
    assert (got.n_minimizers, got.n_anchors, got.n_chains) == (n_mv, n_a, n_chains)
    assert n_regs >= 40 and n_chains >= n_regs

"""Chain -> region post-processing on the device (include/mm2post_b200.h) against the oracle
(oracle/post_oracle.c, itself pinned to the reference's hit.c in tests/test_post_oracle.py), through the C ABI, bit for bit:
every field of every mm_reg1_t and the rewritten anchor arrays."""
import numpy as np
import pytest

from post_util import fuzz_chains, read_hash

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    from minimap2_chaindp_b200 import api
    c = api.ChainContext(device=0, profile=True)
    yield c
    c.close()


def _oracle_regs(preset_post, over, offs, a, res, hashes, qlens):
    from oracle import post as opost
    pp = opost.post_params(preset_post, **over)
    regs, anchors = [], []
    for r in range(offs.shape[0] - 1):
        u = res.u[res.u_off[r]:res.u_off[r + 1]]
        b = res.b[res.b_off[r]:res.b_off[r + 1]]
        g, a2, n_a = opost.chain_post(pp, int(hashes[r]), int(qlens[r]), u, b)
        regs.append(g); anchors.append((a2, n_a))
    return regs, anchors


@pytest.mark.parametrize("chunk", [0, 40000, 1])
def test_batch_runs_as_a_pipeline_of_chunks(ctx, chunk, monkeypatch):
    """MM2B200_REGS_CHUNK anchors per chunk (1 = every read its own chunk; default 8 Mi): uploads of later chunks overlap the
    chaining of earlier ones, regions and anchors of all chunks come back in submission order"""
    import oracle
    from minimap2_chaindp_b200 import api, post, workload
    if chunk:
        monkeypatch.setenv("MM2B200_REGS_CHUNK", str(chunk))
    else:
        monkeypatch.delenv("MM2B200_REGS_CHUNK", raising=False)
    n_reads = 400 if chunk != 1 else 60
    spec = workload.CONFIGS["map-ont-100k"]
    offs, a = workload.generate(spec, 7, n_reads)
    offs = offs + 5                                         # absolute offsets into a larger array
    a = np.concatenate([np.zeros((5, 2), np.uint64), a])
    hashes = np.array([read_hash(r, 3) for r in range(n_reads)], np.uint32)
    qlens = np.full(n_reads, 12000, np.int32)
    want_chains = oracle.chain_batch(oracle.params(spec.preset), offs - 5, a[5:])
    want, want_a = _oracle_regs("map-ont", {}, offs - 5, a[5:], want_chains, hashes, qlens)
    got = post.chain_regs_batch(ctx, api.params(spec.preset), post.post_params("map-ont"), offs, a, hashes, qlens, want_anchors=True)
    assert got.n_chains == int(want_chains.u_off[-1])
    assert np.array_equal(np.diff(got.reg_off), [g.shape[0] for g in want]) and got.reg_off[0] == 0
    assert np.array_equal(got.a_off, want_chains.b_off)
    for r in range(n_reads):
        assert got.regs[got.reg_off[r]:got.reg_off[r + 1]].tobytes() == want[r].tobytes(), f"chunk {chunk} read {r}"
        a2, n_a = want_a[r]
        assert np.array_equal(got.a[got.a_off[r]:got.a_off[r] + n_a], a2[:n_a])


@pytest.mark.parametrize("name,preset_post,n_reads,over", [
    ("map-ont-100k", "map-ont", 3000, {}),
    ("map-ont-100k", "map-ont", 500, dict(pri_ratio=0.3, best_n=1)),
    ("ava-ont-50k", "map-ont", 40, {}),                       # many chains per read through parents / secondaries / joins
    ("ava-ont-50k", "ava-ont", 40, {}),                       # the fork's ava mode: MM_F_ALL_CHAINS, mm_gen_regs only
    ("ava-ont-50k", "map-ont", 40, dict(pri_ratio=0.0)),
    ("map-pb-ul-20k", "map-pb", 12, {}),
])
def test_regions_of_workload_reads_match_the_oracle(ctx, name, preset_post, n_reads, over):
    import oracle
    from minimap2_chaindp_b200 import api, post, workload
    spec = workload.CONFIGS[name]
    offs, a = workload.generate(spec, 0, n_reads)
    hashes = np.array([read_hash(r) for r in range(n_reads)], np.uint32)
    qlens = np.array([int((a[offs[r]:offs[r + 1], 1] & np.uint64(0xffffffff)).max()) + 100 if offs[r + 1] > offs[r] else 100 for r in range(n_reads)], np.int32)
    want_chains = oracle.chain_batch(oracle.params(spec.preset), offs, a)
    want, want_a = _oracle_regs(preset_post, over, offs, a, want_chains, hashes, qlens)
    got = post.chain_regs_batch(ctx, api.params(spec.preset), post.post_params(preset_post, **over), offs, a, hashes, qlens, want_anchors=True)
    assert got.n_chains == int(want_chains.u_off[-1])
    assert np.array_equal(np.diff(got.reg_off), [g.shape[0] for g in want])
    assert np.array_equal(got.a_off, want_chains.b_off)
    for r in range(n_reads):
        g = got.regs[got.reg_off[r]:got.reg_off[r + 1]]
        assert g.tobytes() == want[r].tobytes(), f"{name} read {r}:\n{g}\n{want[r]}"
        a2, n_a = want_a[r]
        assert np.array_equal(got.a[got.a_off[r]:got.a_off[r] + n_a], a2[:n_a]), f"{name} read {r}: anchors"
    assert got.regs.shape[0] > 0
    assert post.post_stage_ms(ctx)[1] >= 3


@pytest.mark.parametrize("seed", range(6))
def test_fuzzed_chain_sets_on_device_resident_chains(ctx, seed):
    """mm2b200_post_device on chains the test places in HBM itself: overlapping query intervals, equal scores, joinable
    neighbours, more than 64 chains per read (the unstable sort of mm_gen_regs), every parameter branch"""
    import torch
    from oracle import post as opost
    from minimap2_chaindp_b200 import api, post
    rng = np.random.default_rng(500 + seed)
    over = [{}, dict(pri_ratio=0.3, best_n=2), dict(mask_level=0.1, min_diff=200), dict(flag=0x400),
            dict(max_join_long=50000, max_join_short=20000, min_join_flank_sc=100), dict(pri_ratio=0.0)][seed]
    us, bs, qlens, hashes = [], [], [], []
    n_reads = 200
    for t in range(n_reads):
        n = 0 if t % 17 == 3 else int(rng.integers(1, 12)) if t % 5 else int(rng.integers(60, 160))
        u, b, qlen = fuzz_chains(rng, n, joinable=0.2 + 0.1 * (seed % 4)) if n else (np.zeros(0, np.uint64), np.zeros((0, 2), np.uint64), 500)
        us.append(u); bs.append(b); qlens.append(qlen); hashes.append(read_hash(t, seed))
    n_u = np.array([u.shape[0] for u in us], np.int32)
    u_off = np.concatenate([[0], np.cumsum(n_u)]).astype(np.int64)
    b_off = np.concatenate([[0], np.cumsum([b.shape[0] for b in bs])]).astype(np.int64)
    dev = torch.device("cuda:0")
    t_nu, t_uoff, t_boff = (torch.from_numpy(x).to(dev) for x in (n_u, u_off, b_off))
    t_u = torch.from_numpy(np.concatenate(us).view(np.int64)).to(dev)
    t_b = torch.from_numpy(np.concatenate(bs).view(np.int64)).to(dev)
    t_hash = torch.from_numpy(np.array(hashes, np.uint32).view(np.int32)).to(dev)
    t_qlen = torch.from_numpy(np.array(qlens, np.int32)).to(dev)
    torch.cuda.synchronize()
    chains = api._DevResult(t_nu.data_ptr(), t_uoff.data_ptr(), t_boff.data_ptr(), t_u.data_ptr(), t_b.data_ptr(), 0)
    out = post.post_device(ctx, post.post_params("map-ont", **over), n_reads, chains, int(u_off[-1]), t_hash.data_ptr(), t_qlen.data_ptr(), 0)
    torch.cuda.synchronize()
    from cuda.bindings import runtime as cudart

    def d2h(ptr, count, dtype):
        t = np.zeros(count, dtype)
        if count:
            (rc,) = cudart.cudaMemcpy(t.ctypes.data, ptr, t.nbytes, cudart.cudaMemcpyKind.cudaMemcpyDeviceToHost)
            assert int(rc) == 0
        return t
    reg_off = d2h(out.d_reg_off, n_reads + 1, np.int64)
    regs = d2h(out.d_regs, int(reg_off[-1]), post.REG_DTYPE)
    b_after = t_b.cpu().numpy().view(np.uint64).reshape(-1, 2)
    pp = opost.post_params("map-ont", **over)
    stats = dict(secondary=0, dropped=0)
    for t in range(n_reads):
        want, a2, n_a = opost.chain_post(pp, hashes[t], qlens[t], us[t], bs[t])
        g = regs[reg_off[t]:reg_off[t + 1]]
        assert g.tobytes() == want.tobytes(), f"seed {seed} read {t} ({us[t].shape[0]} chains):\n{g}\n{want}"
        assert np.array_equal(b_after[b_off[t]:b_off[t] + n_a], a2[:n_a]), f"seed {seed} read {t}: anchors"
        stats["secondary"] += int((want["parent"] != want["id"]).sum()); stats["dropped"] += us[t].shape[0] - want.shape[0]
    assert stats["secondary"] > 0 and (stats["dropped"] > 0 or seed == 5)


def test_empty_and_rejected_inputs(ctx):
    from minimap2_chaindp_b200 import api, post
    pp = post.post_params("map-ont")
    got = post.chain_regs_batch(ctx, api.params("map-ont"), pp, np.zeros(1, np.int64), np.zeros((0, 2), np.uint64), [], [])
    assert got.regs.shape[0] == 0 and got.reg_off.tolist() == [0]
    got = post.chain_regs_batch(ctx, api.params("map-ont"), pp, np.zeros(4, np.int64), np.zeros((0, 2), np.uint64), [1, 2, 3], [10, 10, 10], True)
    assert got.regs.shape[0] == 0 and got.reg_off.tolist() == [0, 0, 0, 0] and got.n_chains == 0
    with pytest.raises(api.ChainError) as ei:                  # multi-segment post-processing is outside this path
        post.chain_regs_batch(ctx, api.params("map-ont", n_segs=2), pp, np.zeros(1, np.int64), np.zeros((0, 2), np.uint64), [], [])
    assert ei.value.code == api.ERR_ARG
    with pytest.raises(api.ChainError):
        post.chain_regs_batch(ctx, api.params("map-ont"), post.post_params("map-ont", max_join_long=0), np.zeros(1, np.int64), np.zeros((0, 2), np.uint64), [], [])

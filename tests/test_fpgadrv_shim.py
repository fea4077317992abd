"""The fork's driver boundary (include/mm2fpgadrv_b200.h) served by the GPU.

1. Whole program: the reference's own minimap2, EVERY object of it (chain.o included), linked against libmm2chain_b200.so in
   place of the closed libfpgadrv.a (oracle/Makefile `dropin` -> oracle/_ref/minimap2_b2), must print the golden PAF lines:
   its request packets (map.c:286-324) are answered by the device.
2. Symbol level: the request packets the unmodified reference built for the fixture runs (tests/golden/seed_*.npz) go through
   fpga_get_writebuf_thread / fpga_writebuf_submit from several threads; what fpga_get_retbuf hands back must hold, read by
   read, the records and mini_pos the reference's own device model produced (map.c:484-575).
3. A read the device refuses comes back with err_flag = 1 and no payload (the fork's protocol, map.c:933-944,970-971) and
   does not fail its neighbours.
"""
import glob
import os
import struct
import subprocess
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "minimap2_b2")
DATA = os.path.join(ROOT, "oracle", "_ref", "test")
GOLD = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "seed_*.npz")))


def _golden_paf():
    out, name = {}, None
    for line in open(os.path.join(ROOT, "tests", "golden", "paf.txt")):
        if line.startswith("## "):
            name = line[3:].split(":")[0]
            out[name] = []
        elif line.strip():
            out[name].append("\t".join(line.rstrip("\n").split("\t")[:18]))
    return out


@pytest.mark.skipif(not os.path.exists(BIN), reason="oracle/_ref/minimap2_b2 not built (needs /root/reference at build time)")
def test_unmodified_reference_program_over_the_driver_boundary(tmp_path):
    both = tmp_path / "both.fa"
    both.write_bytes(open(os.path.join(DATA, "MT-human.fa"), "rb").read() + open(os.path.join(DATA, "MT-orang.fa"), "rb").read())
    runs = {
        "map-ont_MT": ["-x", "map-ont", f"{DATA}/MT-human.fa", f"{DATA}/MT-orang.fa"],
        "default_inv": [f"{DATA}/t-inv.fa", f"{DATA}/q-inv.fa"],
        "default_q2": [f"{DATA}/t2.fa", f"{DATA}/q2.fa"],
        "ava-ont_MT": ["-x", "ava-ont", str(both), str(both)],
    }
    want = _golden_paf()
    for name, args in runs.items():
        p = subprocess.run([BIN, "-t", "12", *args], cwd=tmp_path, capture_output=True, text=True, timeout=300)
        assert p.returncode == 0, p.stderr[-2000:]
        got = ["\t".join(l.split("\t")[:18]) for l in p.stdout.splitlines() if l.strip()]
        assert sorted(got) == sorted(want[name]), f"{name}: PAF differs\n got {got}\nwant {want[name]}\n{p.stderr[-1500:]}"


def _packet(meta, mv, mv_off, reads, magic=0x1234abcd, tid=7):
    """request packet of map.c:286-324 for the given reads"""
    body = b""
    for r in reads:
        gap_qry, gap_ref, qlen, read_id, rank, rflag, n_segs, b = (int(v) for v in meta[r])
        m = mv[mv_off[r]:mv_off[r + 1]]
        body += struct.pack("<iiiiIIhbxQ28x", gap_qry, gap_ref, m.shape[0], qlen, read_id, rank | rflag << 31, n_segs, b, 0)
        raw = m.tobytes()
        body += raw + b"\0" * (-len(raw) % 64)
    return struct.pack("<IIHHBB50x", magic, 64 + len(body), tid, len(reads), 3, 1) + body


def _parse(resp):
    """response packet of map.c:533-565 -> (head fields, {read_id: (err, rec bytes, mini bytes, rep_len)})"""
    magic, size, tid, num, typ, lat = struct.unpack_from("<IIHHBB", resp, 0)
    assert size == len(resp) and typ == 3
    p, out = 64, {}
    for _ in range(num):
        err, read_id, sub, n_a, n_mini, rep_len = struct.unpack_from("<6I", resp, p)
        if err:
            assert sub == 64 and n_a == 0 and n_mini == 0
            out[read_id] = (1, b"", b"", 0)
            p += 64
            continue
        q = p + 64 + (24 * n_a + 63) // 64 * 64
        assert sub == 64 + (24 * n_a + 63) // 64 * 64 + (8 * n_mini + 63) // 64 * 64
        out[read_id] = (0, resp[p + 64:p + 64 + 24 * n_a], resp[q:q + 8 * n_mini], rep_len)
        p += sub
    assert p == len(resp)
    return (magic, tid, num, lat), out


def _pump(packets, n_threads=4):
    """submit from n_threads threads while one receiver collects; returns the responses"""
    from minimap2_chaindp_b200 import fpgadrv
    got, errs = [], []

    def sender(k):
        try:
            for pk in packets[k::n_threads]:
                fpgadrv.submit(pk, tid=k)
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    def receiver():
        while len(got) < len(packets):
            r = fpgadrv.receive()
            if r is None:
                break
            got.append(r)
    ts = [threading.Thread(target=sender, args=(k,)) for k in range(n_threads)] + [threading.Thread(target=receiver)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=300)
    assert not errs and len(got) == len(packets)
    return got


@pytest.mark.parametrize("zerocopy", ["1", "0"])
def test_request_packets_of_the_reference_through_the_driver_symbols(monkeypatch, zerocopy):
    """both data paths of the shim: packets DMA'd as they are and laid out on the device (default), and the careful path that
    stages and cuts them on the host (MM2B200_SHIM_ZEROCOPY=0; also what a refused batch falls back to)"""
    from minimap2_chaindp_b200 import fpgadrv
    monkeypatch.setenv("MM2B200_SHIM_ZEROCOPY", zerocopy)
    fpgadrv.init()
    try:
        for path in GOLD:                                   # each fixture loads a new index over the old one
            z = np.load(path)
            fpgadrv.load_index(z["idx_B"], fpgadrv.TYPE_INDEX_B, chunk=100000)
            fpgadrv.load_index(z["idx_H"], fpgadrv.TYPE_INDEX_H, chunk=65536)
            fpgadrv.load_index(z["idx_V"], fpgadrv.TYPE_INDEX_V)
            fpgadrv.load_index(z["idx_P"], fpgadrv.TYPE_INDEX_P)
            fpgadrv.set_params(*(int(v) for v in z["params"]))
            meta, n = z["meta"], int(z["n_reads"][0])
            packets = [_packet(meta, z["mv"], z["mv_off"], range(r0, min(n, r0 + 8)), magic=1000 + r0, tid=r0 % 5) for r0 in range(0, n, 8)]
            before = fpgadrv.stats()
            got = _pump(packets)
            seen = {}
            for resp in got:
                (magic, tid, num, lat), reads = _parse(resp)
                r0 = magic - 1000
                assert tid == r0 % 5 and num == min(n, r0 + 8) - r0 and lat == 1
                assert sorted(reads) == sorted(int(meta[r, 3]) for r in range(r0, r0 + num))
                seen.update(reads)
            assert len(seen) == n
            for r in range(n):
                err, rec, mini, rep_len = seen[int(meta[r, 3])]
                assert err == 0 and rep_len == int(z["rep_len"][r])
                assert rec == z["rec"][24 * z["rec_off"][r]:24 * z["rec_off"][r + 1]].tobytes(), f"{os.path.basename(path)}: records of read {r}"
                assert mini == z["mini"][z["mini_off"][r]:z["mini_off"][r + 1]].tobytes()
            after = fpgadrv.stats()
            assert after["packets"] - before["packets"] == len(packets) and after["reads"] - before["reads"] == n
            assert after["launches"] > before["launches"] and after["err_reads"] == before["err_reads"]
            assert 1 <= after["batches"] - before["batches"] <= len(packets)       # packets are combined into device batches
        # the receiver's exit signal (fpga_chaindp.c:242-245), and re-arming
        fpgadrv.exit_block()
        assert fpgadrv.receive() is None
        fpgadrv.set_block()
    finally:
        fpgadrv.finalize()


def test_a_read_the_device_refuses_goes_back_with_err_flag():
    """one minimizer with 3000 occurrences repeated 800 times in a read: 2.4 M anchors, beyond the 2^21 per-read index space
    of the seeding stage; its neighbours in the same packet are answered normally"""
    from oracle import seed as oseed
    from minimap2_chaindp_b200 import api, fpgadrv, seed
    rng = np.random.default_rng(21)
    table = {}
    while len(table) < 500:
        table[int(rng.integers(1 << 34))] = [oseed.pack_position(int(rng.integers(4)), int(rng.integers(1 << 16)), int(rng.integers(2)), int(rng.integers(6)))
                                               for _ in range(int(rng.integers(1, 4)))]
    heavy = 12345678901
    table[heavy] = [oseed.pack_position(k % 4, 7 * k + 11, k & 1, k % 6) for k in range(3000)]
    oix = oseed.build_streams(table, b=6, rng=rng, load=0.7, deleted=0.1)
    keys = np.array([k for k in table if k != heavy], np.uint64)

    def read(n, key=None):
        ks = keys[rng.integers(0, keys.size, n)] if key is None else np.full(n, key, np.uint64)
        pos = np.cumsum(rng.integers(1, 12, n)) + 20
        m = np.zeros((n, 2), np.uint64)
        m[:, 0] = (ks << np.uint64(8)) | np.uint64(15)
        m[:, 1] = pos.astype(np.uint64) << np.uint64(1)
        return m, int(pos[-1]) + 40
    parts = [read(200), read(800, heavy), read(300)]
    mv = np.concatenate([p[0] for p in parts])
    mv_off = np.array([0, 200, 1000, 1300], np.int64)
    meta = np.array([[5000, 5000, parts[r][1], 40 + r, r, 0, 1, 6] for r in range(3)], np.int64)
    max_occ = 5000
    fpgadrv.init()
    try:
        for s, t in ((oix.B, fpgadrv.TYPE_INDEX_B), (oix.H, fpgadrv.TYPE_INDEX_H), (oix.V, fpgadrv.TYPE_INDEX_V), (oix.P, fpgadrv.TYPE_INDEX_P)):
            fpgadrv.load_index(s, t)
        fpgadrv.set_params(500, 0, 25, 40, 0, max_occ)
        fpgadrv.submit(_packet(meta, mv, mv_off, range(3)))
        _, reads = _parse(fpgadrv.receive())
        assert fpgadrv.stats()["err_reads"] == 1
    finally:
        fpgadrv.finalize()
    assert reads[41] == (1, b"", b"", 0)
    # the neighbours: the same device path called directly (itself pinned to the oracle in test_gpu_seed.py)
    with api.ChainContext(device=0) as ctx:
        ix = seed.SeedIndex(ctx).load_streams(oix.B, oix.H, oix.V, oix.P)
        prm = seed.seed_params(api.params("map-ont", max_dist_x=5000, max_dist_y=5000, bw=500, max_skip=25, min_sc=40, n_segs=5000), 0, max_occ)
        for r in (0, 2):
            out = ix.seed_fill_batch(prm, np.array([0, mv_off[r + 1] - mv_off[r]], np.int64), mv[mv_off[r]:mv_off[r + 1]], [r], [parts[r][1]])
            err, rec, mini, rep_len = reads[40 + r]
            assert err == 0 and rec == out.rec.tobytes() and len(rec) > 0 and mini == out.mini_pos.tobytes() and rep_len == int(out.rep_len[0])
            a, rep, mini_o = oseed.collect_seed_hits(oix, 0, max_occ, mv[mv_off[r]:mv_off[r + 1]], r, parts[r][1])
            assert mini == mini_o.tobytes() and rep_len == rep
        ix.close()


def test_both_data_paths_answer_with_the_same_bytes(monkeypatch):
    """ragged packets (a read without minimizers, a packet without reads, packets of 1..8 reads) through the direct path and
    the careful path: every response packet must be byte-identical, padding included"""
    from minimap2_chaindp_b200 import fpgadrv
    z = np.load(GOLD[0])
    meta, n, mv, mv_off = z["meta"].copy(), int(z["n_reads"][0]), z["mv"], z["mv_off"].copy()
    rng = np.random.default_rng(5)
    # read 1 loses its minimizers (seednum = 0): shift nothing, just give it an empty slice through a private offset table
    lo, hi = mv_off[:-1].copy(), mv_off[1:].copy()
    hi[1] = lo[1]

    def packet(reads, magic):
        body = b""
        for r in reads:
            gap_qry, gap_ref, qlen, read_id, rank, rflag, n_segs, b = (int(v) for v in meta[r])
            m = mv[lo[r]:hi[r]]
            body += struct.pack("<iiiiIIhbxQ28x", gap_qry, gap_ref, m.shape[0], qlen, read_id, rank | rflag << 31, n_segs, b, 0)
            raw = m.tobytes()
            body += raw + b"\0" * (-len(raw) % 64)
        return struct.pack("<IIHHBB50x", magic, 64 + len(body), magic % 3, len(reads), 3, 1) + body
    packets, r0, magic = [], 0, 500
    while r0 < n:
        k = int(rng.integers(1, 9))
        packets.append(packet(range(r0, min(n, r0 + k)), magic))
        r0 += k
        magic += 1
        if magic % 7 == 0:
            packets.append(packet([], magic))              # a packet without reads
            magic += 1
    answers = {}
    for mode in ("1", "0"):
        monkeypatch.setenv("MM2B200_SHIM_ZEROCOPY", mode)
        fpgadrv.init()
        try:
            fpgadrv.load_index(z["idx_B"], fpgadrv.TYPE_INDEX_B)
            fpgadrv.load_index(z["idx_H"], fpgadrv.TYPE_INDEX_H)
            fpgadrv.load_index(z["idx_V"], fpgadrv.TYPE_INDEX_V)
            fpgadrv.load_index(z["idx_P"], fpgadrv.TYPE_INDEX_P)
            fpgadrv.set_params(*(int(v) for v in z["params"]))
            got = _pump(packets)
            assert fpgadrv.stats()["err_reads"] == 0
        finally:
            fpgadrv.finalize()
        answers[mode] = {struct.unpack_from("<I", g, 0)[0]: bytes(g) for g in got}
        assert len(answers[mode]) == len(packets)
    for mg, resp in answers["1"].items():
        assert resp == answers["0"][mg], f"response to packet {mg} differs between the two paths"
        _parse(resp)
    # and against the golden records where the read kept its minimizers
    seen = {}
    for resp in answers["1"].values():
        seen.update(_parse(resp)[1])
    for r in range(n):
        err, rec, mini, rep_len = seen[int(meta[r, 3])]
        if r == 1:
            assert (err, rec, mini) == (0, b"", b"")
            continue
        assert err == 0 and rec == z["rec"][24 * z["rec_off"][r]:24 * z["rec_off"][r + 1]].tobytes() and rep_len == int(z["rep_len"][r])

#!/usr/bin/env python
"""pack_seed_golden.py <dump> <out.npz> -- turns the records loopback_stub.c wrote ($MM2_SEED_DUMP) into one fixture:
index streams, device parameters, per-read minimizers + task header fields, and the anchors the reference's own
collect_seed_hits produced for each read (in packet order)."""
import struct
import sys

import numpy as np

HDR = 64      # sizeof(chaindp_sndhdr_t), fpga_chaindp.h:79-87
TASK = 64     # sizeof(collect_task_t), fpga_chaindp.h:46-58
RES = 64      # sizeof(collect_result_t), fpga_chaindp.h:60-68


def main(dump, out):
    raw = open(dump, "rb").read()
    pos = 0
    params = None
    idx = {4: [], 5: [], 6: [], 7: []}
    reads, anchors, results = [], [], []
    while pos < len(raw):
        tag = raw[pos:pos + 1]; pos += 1
        if tag == b"P":
            params = struct.unpack_from("<6i", raw, pos); pos += 24
        elif tag == b"I":
            t, size = struct.unpack_from("<2i", raw, pos); pos += 8
            idx[t].append(raw[pos:pos + size]); pos += size
        elif tag == b"K":
            (size,) = struct.unpack_from("<i", raw, pos); pos += 4
            pk = raw[pos:pos + size]; pos += size
            magic, psize, tid, num, typ, lat = struct.unpack_from("<IIHHBB", pk, 0)
            p = HDR
            for _ in range(num):
                gap_qry, gap_ref, seednum, qlensum, read_id, bid, n_segs, b = struct.unpack_from("<iiiiIIhb", pk, p)
                mv = np.frombuffer(pk, np.uint64, 2 * seednum, p + TASK).reshape(-1, 2).copy()
                reads.append(dict(gap_qry=gap_qry, gap_ref=gap_ref, qlen=qlensum, read_id=read_id, bid=bid, n_segs=n_segs, b=b, mv=mv))
                p += TASK + (seednum * 16 + 63) // 64 * 64
        elif tag == b"R":
            (size,) = struct.unpack_from("<i", raw, pos); pos += 4
            pk = raw[pos:pos + size]; pos += size
            magic, psize, tid, num, typ, lat = struct.unpack_from("<IIHHBB", pk, 0)
            p = HDR
            for _ in range(num):
                err, read_id, sub_size, n_a, n_mini, rep_len = struct.unpack_from("<6I", pk, p)
                rec = np.frombuffer(pk, np.uint8, 24 * n_a, p + RES).copy()      # struct new_seed[n_a], minimap.h:51-55
                q = p + RES + (24 * n_a + 63) // 64 * 64
                mini = np.frombuffer(pk, np.uint64, n_mini, q).copy()
                results.append(dict(err=err, read_id=read_id, rep_len=rep_len, rec=rec, mini=mini))
                assert sub_size == RES + (24 * n_a + 63) // 64 * 64 + (8 * n_mini + 63) // 64 * 64
                p += sub_size
        elif tag == b"A":
            (n,) = struct.unpack_from("<q", raw, pos); pos += 8
            anchors.append(np.frombuffer(raw, np.uint64, 2 * n, pos).reshape(-1, 2).copy()); pos += 16 * n
        else:
            raise SystemExit(f"bad tag {tag!r} at {pos}")
    assert params is not None and len(reads) == len(anchors) == len(results), (len(reads), len(anchors), len(results))
    assert all(r["read_id"] == q["read_id"] and q["err"] == 0 for r, q in zip(reads, results))
    d = {"params": np.array(params, np.int32)}     # bw, is_cdna, max_skip, min_sc, flag, max_occ
    for t, name in ((4, "idx_B"), (5, "idx_H"), (6, "idx_V"), (7, "idx_P")):
        d[name] = np.frombuffer(b"".join(idx[t]), np.uint8).copy()
    d["n_reads"] = np.array([len(reads)], np.int32)
    d["meta"] = np.array([[r["gap_qry"], r["gap_ref"], r["qlen"], r["read_id"], r["bid"] & 0x7fffffff, (r["bid"] >> 31) & 1, r["n_segs"], r["b"]]
                          for r in reads], np.int64)    # gap_qry, gap_ref, qlen, read_id, bid rank, bid flag, n_segs, b
    d["mv_off"] = np.concatenate([[0], np.cumsum([r["mv"].shape[0] for r in reads])]).astype(np.int64)
    d["mv"] = np.concatenate([r["mv"] for r in reads]) if reads else np.zeros((0, 2), np.uint64)
    d["a_off"] = np.concatenate([[0], np.cumsum([a.shape[0] for a in anchors])]).astype(np.int64)
    d["a"] = np.concatenate(anchors) if anchors else np.zeros((0, 2), np.uint64)
    d["rep_len"] = np.array([q["rep_len"] for q in results], np.int64)
    d["rec_off"] = np.concatenate([[0], np.cumsum([len(q["rec"]) // 24 for q in results])]).astype(np.int64)
    d["rec"] = np.concatenate([q["rec"] for q in results]) if results else np.zeros(0, np.uint8)     # raw new_seed bytes
    d["mini_off"] = np.concatenate([[0], np.cumsum([len(q["mini"]) for q in results])]).astype(np.int64)
    d["mini"] = np.concatenate([q["mini"] for q in results]) if results else np.zeros(0, np.uint64)
    np.savez_compressed(out, **d)
    print(f"{out}: {len(reads)} reads, {int(d['mv_off'][-1])} minimizers, {int(d['a_off'][-1])} anchors, "
          f"index B/H/V/P = {[len(d[k]) for k in ('idx_B', 'idx_H', 'idx_V', 'idx_P')]} bytes, params {params}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])

#!/usr/bin/env python
"""pack_golden.py <dump> <out.npz> -- joins the 'F' (mm_chain_dp_fpga) and 'B' (mm_chain_dp_bottom) records that
dump_hook.c wrote during one reference run into one golden file: per call, the anchors, every DP parameter, the
new_seed records and the final (u, b)."""
import struct
import sys

import numpy as np

SEED = np.dtype([("x", "<u8"), ("y", "<u8"), ("p", "<i4"), ("f", "<i4")])


def main(dump, out):
    try:
        raw = open(dump, "rb").read()
    except FileNotFoundError:
        raw = b""
    pos, F, B = 0, [], []
    while pos < len(raw):
        tag = raw[pos:pos + 1]
        pos += 1
        if tag == b"F":
            prm = struct.unpack_from("<8i", raw, pos); pos += 32
            (n,) = struct.unpack_from("<q", raw, pos); pos += 8
            a = np.frombuffer(raw, np.uint64, 2 * n, pos).reshape(-1, 2).copy(); pos += 16 * n
            (m,) = struct.unpack_from("<I", raw, pos); pos += 4
            rec = np.frombuffer(raw, SEED, m, pos).copy(); pos += 24 * m
            F.append((prm, a, rec))
        elif tag == b"B":
            prm = struct.unpack_from("<3i", raw, pos); pos += 12
            (m,) = struct.unpack_from("<I", raw, pos); pos += 4
            rec = np.frombuffer(raw, SEED, m, pos).copy(); pos += 24 * m
            (nu,) = struct.unpack_from("<i", raw, pos); pos += 4
            u = np.frombuffer(raw, np.uint64, nu, pos).copy(); pos += 8 * nu
            (nb,) = struct.unpack_from("<q", raw, pos); pos += 8
            b = np.frombuffer(raw, np.uint64, 2 * nb, pos).reshape(-1, 2).copy(); pos += 16 * nb
            B.append((prm, rec, u, b))
        else:
            raise SystemExit(f"bad tag {tag!r} at {pos}")
    # join on identical record arrays (the response packet carries a copy of new_seed[], map.c:527-540)
    used = set()
    calls = {}
    F.sort(key=lambda t: (t[1].shape[0], t[1].tobytes()))   # deterministic order (threads finish in any order)
    for k, (fp, a, rec) in enumerate(F):
        match = None
        for j, (bp, brec, u, b) in enumerate(B):
            if j not in used and brec.shape == rec.shape and brec.tobytes() == rec.tobytes():
                match = j
                break
        if match is None:   # should not happen: every response packet goes through mm_chain_dp_bottom (map.c:837)
            raise SystemExit(f"call {k} (n={a.shape[0]}) has no matching mm_chain_dp_bottom record")
        used.add(match)
        bp, _, u, b = B[match]
        # max_dist_x, max_dist_y, bw, max_skip, min_cnt, min_sc, is_cdna, n_segs (argument order of mm_chain_dp, chain.c:23)
        assert bp[1] == fp[4], "min_sc differs between the two halves"
        calls[f"c{k}_params"] = np.array([fp[0], fp[1], fp[2], fp[3], bp[0], fp[4], fp[5], fp[6]], np.int32)
        calls[f"c{k}_n_segs_bottom"] = np.array([bp[2]], np.int32)
        calls[f"c{k}_a"] = a
        calls[f"c{k}_rec"] = rec
        calls[f"c{k}_u"] = u
        calls[f"c{k}_b"] = b
    calls["n_calls"] = np.array([len(F)], np.int32)
    np.savez_compressed(out, **calls)
    print(f"{out}: {len(F)} DP calls, n = {[int(f[1].shape[0]) for f in F]}, "
          f"chains = {[[(int(x >> 32), int(x & 0xffffffff)) for x in calls[f'c{k}_u']] for k in range(len(F))]}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])

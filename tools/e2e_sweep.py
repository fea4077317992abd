#!/usr/bin/env python
"""End-to-end (pinned host anchors in, pinned host chains out) step time for several sub-batch sizes / slot counts.

    python tools/e2e_sweep.py [--workload map-ont-100k] [--reads 100000] [--reps 5]
"""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="map-ont-100k")
    ap.add_argument("--reads", type=int, default=100000)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--grid", default="3:8,3:4,4:4,4:2,6:2,3:16,2:8")   # n_slots:slot_anchors(Mi)
    args = ap.parse_args()
    import torch
    from minimap2_chaindp_b200 import api, workload
    spec = workload.CONFIGS[args.workload]
    offs = workload.count(spec, 0, args.reads)
    pin = api.PinnedArray((int(offs[-1]), 2), np.uint64)
    workload.fill(spec, offs, pin.array, 0)
    prm = api.params(spec.preset)
    print(f"{args.workload} reads={args.reads} anchors={int(offs[-1])} H2D={int(offs[-1]) * 16 / 1e9:.3f} GB")
    for cell in args.grid.split(","):
        ns, sa = (int(x) for x in cell.split(":"))
        with api.ChainContext(device=0, n_slots=ns, slot_anchors=sa << 20) as ctx:
            for _ in range(2):
                res = ctx.chain_batch_raw(prm, args.reads, offs.ctypes.data, pin.ptr)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(args.reps):
                res = ctx.chain_batch_raw(prm, args.reads, offs.ctypes.data, pin.ptr)
            torch.cuda.synchronize()
            dt = (time.perf_counter() - t0) / args.reps
            print(f"slots={ns} slot_anchors={sa}Mi sub_batches={int(res.n_sub_batches)} e2e_ms={dt * 1e3:.2f}", flush=True)


if __name__ == "__main__":
    main()

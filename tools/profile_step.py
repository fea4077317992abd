#!/usr/bin/env python
"""Minimal driver for ncu: a few device-resident passes of the hot path on a slice of a workload.

    python tools/profile_step.py [--workload map-ont-100k] [--reads 20000] [--steps 2]
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
    ap.add_argument("--reads", type=int, default=20000)
    ap.add_argument("--first", type=int, default=0)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--e2e", action="store_true")
    args = ap.parse_args()
    import torch
    from minimap2_chaindp_b200 import api, workload
    spec = workload.CONFIGS[args.workload]
    offs, a = workload.generate(spec, args.first, args.reads)
    prm = api.params(spec.preset)
    dev = torch.device("cuda", 0)
    d_a = torch.from_numpy(a.view(np.int64)).to(dev)
    d_offs = torch.from_numpy(offs).to(dev)
    ctx = api.ChainContext(device=0, profile=True)
    st = torch.cuda.current_stream()
    for k in range(args.steps + 1):
        if k == 1:
            torch.cuda.synchronize()
            ctx.stage_reset()
            t0 = time.perf_counter()
        ctx.chain_device(prm, args.reads, int(offs[-1]), d_offs.data_ptr(), d_a.data_ptr(), st.cuda_stream)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / args.steps
    stages = {s: round(ctx.stage_ms(s)[0] / args.steps, 3) for s in ("prep", "fill", "compact", "tail", "gather")}
    print(f"{args.workload} reads={args.reads} anchors={int(offs[-1])} ms/step={dt * 1e3:.3f} stages={stages}")
    if args.e2e:
        res = ctx.chain_batch(prm, offs, a)
        print("e2e chains:", int(res.u_off[-1]), "visited:", res.n_visited)
    ctx.close()


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Can the HBM-bound stages of one sub-batch run under the issue-bound fill of another?

Two contexts on one GPU, each with half of the reads resident in HBM, driven from two host threads on two streams,
against one context with all reads.  MM2B200_FILL_CTAS caps the resident fill CTAs per SM so that the streaming kernels
of the other half find room.

    MM2B200_FILL_CTAS=8 python tools/overlap_probe.py [--reads 100000] [--steps 10]
"""
import argparse
import os
import sys
import threading
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="map-ont-100k")
    ap.add_argument("--reads", type=int, default=100000)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--parts", type=int, default=2)
    args = ap.parse_args()
    import torch
    from minimap2_chaindp_b200 import api, workload
    spec = workload.CONFIGS[args.workload]
    prm = api.params(spec.preset)
    dev = torch.device("cuda", 0)

    def resident(first, n):
        offs, a = workload.generate(spec, first, n)
        return n, int(offs[-1]), torch.from_numpy(offs).to(dev), torch.from_numpy(a.view(np.int64)).to(dev)

    def timed(jobs):
        """jobs: list of (ctx, stream, n, na, d_offs, d_a); each job is enqueued from its own host thread"""
        def run(job, steps):
            ctx, st, n, na, d_offs, d_a = job
            for _ in range(steps):
                ctx.chain_device(prm, n, na, d_offs.data_ptr(), d_a.data_ptr(), st.cuda_stream)
        for steps in (2, args.steps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            th = [threading.Thread(target=run, args=(j, steps)) for j in jobs]
            [t.start() for t in th]
            [t.join() for t in th]
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
        return 1e3 * dt / args.steps

    whole = resident(0, args.reads)
    ctx = api.ChainContext(device=0)
    one = timed([(ctx, torch.cuda.Stream(), *whole)])
    per = args.reads // args.parts
    jobs = [(api.ChainContext(device=0), torch.cuda.Stream(), *resident(k * per, per)) for k in range(args.parts)]
    split = timed(jobs)
    print(f"{spec.name} reads={args.reads} fill_ctas_cap={os.environ.get('MM2B200_FILL_CTAS', '0')}: "
          f"one context {one:.3f} ms/step, {args.parts} contexts on {args.parts} streams {split:.3f} ms/step")


if __name__ == "__main__":
    main()

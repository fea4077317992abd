#!/usr/bin/env python
"""Minimal driver for ncu: a few device-resident passes of the sketching stage (mm2b200_sketch_device) on G-seq reads.

    python tools/profile_sketch.py [--reads 20000] [--steps 2] [--preset map-ont]
"""
import argparse
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=20000)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--preset", default="map-ont")
    args = ap.parse_args()
    import torch
    from minimap2_chaindp_b200 import api, sketch, workload
    W = workload.SeqWorkload(workload.SEQ_CONFIGS["map-ont-small-seq"])
    off, _ = W.reads_count(0, args.reads)
    seq = np.zeros(int(off[-1]), np.uint8)
    W.reads_fill(0, off, seq)
    w, k, hpc = sketch.SKETCH_PRESETS[args.preset]
    dev = torch.device("cuda", 0)
    d_seq, d_off = torch.from_numpy(seq).to(dev), torch.from_numpy(off).to(dev)
    prm = sketch.SketchParams(w, k, hpc)
    out, n = sketch._DevSketch(), C.c_int64()
    with api.ChainContext(device=0) as ctx:
        for it in range(args.steps + 1):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            rc = sketch._lib().mm2b200_sketch_device(ctx._h, C.byref(prm), args.reads, int(off[-1]), d_off.data_ptr(), d_seq.data_ptr(), None, None,
                                                     C.byref(out), C.byref(n))
            assert rc == 0, api.last_error()
            torch.cuda.synchronize()
            print(f"pass {it}: {1e3 * (time.perf_counter() - t0):.3f} ms, {int(off[-1])} bases, {n.value} minimizers")


if __name__ == "__main__":
    main()

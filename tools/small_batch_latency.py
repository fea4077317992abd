#!/usr/bin/env python
"""Latency of mm2b200_chain_batch for small batches (what the level-1 dispatcher pays per device batch)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from minimap2_chaindp_b200 import api, workload
spec = workload.CONFIGS["map-ont-100k"]
prm = api.params(spec.preset)
with api.ChainContext(device=0, n_slots=2, slot_anchors=4 << 20) as c:
    for nr in (1, 8, 64, 512, 4096):
        offs = workload.count(spec, 0, nr)
        pa = api.PinnedArray((int(offs[-1]), 2), np.uint64)
        workload.fill(spec, offs, pa.array, 0)
        for _ in range(20): c.chain_batch_raw(prm, nr, offs.ctypes.data, pa.ptr)
        t0 = time.perf_counter(); n = 200
        for _ in range(n): c.chain_batch_raw(prm, nr, offs.ctypes.data, pa.ptr)
        dt = (time.perf_counter() - t0) / n
        print(f"{nr:5d} reads/batch: {dt*1e6:8.1f} us per batch, {nr/dt:10.0f} reads/s")

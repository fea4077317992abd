#!/usr/bin/env python
"""Small end-to-end case for compute-sanitizer (memcheck / racecheck): every kernel variant, a few hundred reads."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from minimap2_chaindp_b200 import api, workload  # noqa: E402
from util import concat_reads, many_chain_reads, random_colinear, random_dense  # noqa: E402

rng = np.random.default_rng(3)
cases = []
offs, a = workload.generate(workload.CONFIGS["map-ont-100k"], 0, 150)
cases.append(("map-ont", {}, offs, a))
offs, a = workload.generate(workload.CONFIGS["ava-ont-50k"], 0, 6)
cases.append(("ava-ont", {}, offs, a))
offs, a = concat_reads([random_colinear(rng, int(n), noise=0.5, step=3) for n in rng.integers(2, 1500, 12)])
cases.append(("map-ont", dict(max_skip=3), offs, a))
cases.append(("map-ont", dict(bw=2000), offs, a))              # no gap-cost table
cases.append(("map-ont", dict(max_iter=40), offs, a))           # FULLV variant
offs, a = concat_reads([random_dense(rng, int(n), segs=2) for n in rng.integers(0, 200, 30)])
cases.append(("sr", {}, offs, a))                               # generic kernel
cases.append(("splice", {}, offs, a))
offs, a = concat_reads(many_chain_reads(rng, 2, 120))
cases.append(("map-ont", dict(min_sc=10, min_cnt=2), offs, a))  # > 64 chains: exact radix replay
with api.ChainContext(device=0, n_slots=2, slot_anchors=40000, count_visited=True) as c:
    for preset, over, offs, a in cases:
        r = c.chain_batch(api.params(preset, **over), offs, a)
        ro, rec = c.fill_batch(api.params(preset, **over), offs, a)
        r2 = c.bottom_batch(api.params(preset, **over), ro, rec)
        assert np.array_equal(r.u, r2.u) and np.array_equal(r.b, r2.b)
        print(preset, over, "chains", int(r.u_off[-1]), "visited", r.n_visited)
print("sanitize case done")

#!/usr/bin/env python
"""Level-1 drop-in throughput: T host threads each calling the reference entry point mm_chain_dp one read at a time
(what the kt_for workers of map.c do); the library's dispatcher combines concurrent callers into device batches."""
import os
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from minimap2_chaindp_b200 import api, workload  # noqa: E402

spec = workload.CONFIGS["map-ont-100k"]
offs, a = workload.generate(spec, 0, 16384 if "--dump" in sys.argv else 4096)
prm = api.params(spec.preset)
if "--dump" in sys.argv:                           # input file for the C harness (tools/dropin_rate.c)
    import numpy as np
    with open(sys.argv[sys.argv.index("--dump") + 1], "wb") as fh:
        np.array([offs.shape[0] - 1], np.int64).tofile(fh); offs.tofile(fh); a.tofile(fh)
    sys.exit(0)
api.mm_chain_dp(prm, a[offs[0]:offs[1]])          # creates the default context
for T in (1, 16, 64, 256):
    n_per = 4096 // T
    def work(t):
        for r in range(t * n_per, (t + 1) * n_per):
            api.mm_chain_dp(prm, a[offs[r]:offs[r + 1]])
    th = [threading.Thread(target=work, args=(t,)) for t in range(T)]
    t0 = time.perf_counter()
    [x.start() for x in th]
    [x.join() for x in th]
    dt = time.perf_counter() - t0
    print(f"{T:4d} caller threads: {T * n_per / dt:9.0f} reads/s through mm_chain_dp (python ctypes callers)")

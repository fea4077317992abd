"""Single-process multi-GPU sharding through mm2b200_chain_batch_multi (INTEGRATION.md level 2)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run(ctxs, n_reads=2500):
    import oracle
    from minimap2_chaindp_b200 import api, workload
    from util import assert_same
    spec = workload.CONFIGS["map-ont-100k"]
    offs, a = workload.generate(spec, 0, n_reads)
    want = oracle.chain_batch(oracle.params(spec.preset), offs, a)
    got, cuts = api.chain_batch_multi(ctxs, api.params(spec.preset), offs, a)
    assert cuts[0] == 0 and cuts[-1] == n_reads
    assert_same(got, want, "multi")
    return cuts


def test_chain_batch_multi_on_distinct_gpus():
    """One context per PHYSICAL GPU (skipped, not aliased, on a 1-GPU box)."""
    from minimap2_chaindp_b200 import api
    n_dev = api.lib().mm2b200_device_count()
    if n_dev < 2:
        pytest.skip(f"{n_dev} GPU visible: the multi-GPU path needs at least 2")
    n = min(n_dev, 8)
    ctxs = [api.ChainContext(device=k, slot_anchors=300000) for k in range(n)]
    try:
        assert len({c.device for c in ctxs}) == n
        cuts = _run(ctxs, 4000)
        assert np.all(np.diff(cuts) > 0), "every GPU must get a share"
    finally:
        for c in ctxs:
            c.close()


def test_chain_batch_multi_threaded_path_one_gpu():
    """The host-thread-per-context machinery with two contexts on device 0 (what a 1-GPU box can check)."""
    from minimap2_chaindp_b200 import api
    ctxs = [api.ChainContext(device=0, slot_anchors=300000) for _ in range(2)]
    try:
        _run(ctxs)
    finally:
        for c in ctxs:
            c.close()


def test_driver_boundary_shares_packets_among_gpus():
    """The driver shim with one worker per GPU (MM2B200_DEVICES=all): the index is replicated, the packets of the G-seq fixture
    are answered by whichever device is free, every read's records equal the reference's (run in a subprocess: the shim reads
    its device list at fpga_init)."""
    import os
    import subprocess
    import sys
    from minimap2_chaindp_b200 import api
    n_dev = api.lib().mm2b200_device_count()
    if n_dev < 2:
        pytest.skip(f"{n_dev} GPU visible: the multi-GPU path needs at least 2")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = r'''
import sys, numpy as np
sys.path.insert(0, "tests")
import test_fpgadrv_shim as T
from minimap2_chaindp_b200 import fpgadrv
z = np.load("tests/golden/seed_map-ont_gseq.npz")
fpgadrv.init()
for key, t in (("idx_B", 4), ("idx_H", 5), ("idx_V", 6), ("idx_P", 7)):
    fpgadrv.load_index(z[key], t)
fpgadrv.set_params(*(int(v) for v in z["params"]))
meta, n = z["meta"], int(z["n_reads"][0])
seen = {}
for rep in range(6):                      # several rounds so that every worker gets packets
    packets = [T._packet(meta, z["mv"], z["mv_off"], range(r0, min(n, r0 + 4)), magic=r0) for r0 in range(0, n, 4)]
    for resp in T._pump(packets, n_threads=8):
        seen.update(T._parse(resp)[1])
for r in range(n):
    err, rec, mini, rep_len = seen[int(meta[r, 3])]
    assert err == 0 and rec == z["rec"][24 * z["rec_off"][r]:24 * z["rec_off"][r + 1]].tobytes()
st = fpgadrv.stats()
fpgadrv.finalize()
assert st["err_reads"] == 0 and st["batches"] >= 6
print("ok", st)
'''
    env = dict(os.environ, MM2B200_DEVICES="all")
    p = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0 and "ok" in p.stdout, p.stdout[-500:] + p.stderr[-2000:]

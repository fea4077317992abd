"""Single-process multi-GPU sharding through mm2b200_chain_batch_multi (needs >= 2 GPUs; skipped otherwise)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_chain_batch_multi_matches_oracle():
    import oracle
    from minimap2_chaindp_b200 import api, workload
    from util import assert_same
    n_dev = api.lib().mm2b200_device_count()
    n = min(n_dev, 4)
    spec = workload.CONFIGS["map-ont-100k"]
    offs, a = workload.generate(spec, 0, 2500)
    want = oracle.chain_batch(oracle.params(spec.preset), offs, a)
    # with one GPU this still exercises the threaded path: two contexts on the same device
    ctxs = [api.ChainContext(device=k % n_dev, slot_anchors=300000) for k in range(max(n, 2))]
    try:
        got, cuts = api.chain_batch_multi(ctxs, api.params(spec.preset), offs, a)
        assert cuts[0] == 0 and cuts[-1] == 2500
        assert_same(got, want, "multi")
    finally:
        for c in ctxs:
            c.close()

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

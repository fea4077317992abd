"""Stress of the fill kernel's intra-warp ordering assumptions (ring stores by all lanes, marks through tag words / t[] +
__syncwarp): the same adversarial inputs under every residency (MM2B200_FILL_CTAS = 1 .. max), ring depth and slot count,
many repetitions, digests compared with the oracle's."""
import hashlib

import numpy as np
import pytest

import oracle
from minimap2_chaindp_b200 import api
from util import concat_reads, interleaved_diagonals, random_colinear

pytestmark = pytest.mark.gpu


def _digest(res):
    h = hashlib.sha1()
    for arr in (res.n_u, res.u_off, res.b_off, res.u, res.b):
        h.update(np.ascontiguousarray(arr).tobytes())
    return h.hexdigest()


def test_fill_is_deterministic_under_every_residency(monkeypatch):
    rng = np.random.default_rng(99)
    reads = [interleaved_diagonals(rng, int(n), k, step=int(rng.integers(1, 4))) for k in (2, 5, 20, 50) for n in rng.integers(300, 3000, 6)]
    reads += [random_colinear(rng, int(n), noise=float(rng.random()) * 0.6, step=int(rng.integers(1, 8))) for n in rng.integers(2, 3000, 30)]
    offs, a = concat_reads(reads)
    prm, oprm = api.params("map-ont"), oracle.params("map-ont")
    want = _digest(oracle.chain_batch(oprm, offs, a))
    runs = 0
    for ctas in (1, 2, 5, 0):
        for ring in (64, -1, 256):
            for slots, slot_anchors in ((1, 1 << 14), (2, 1 << 12), (32, 1 << 11)):
                monkeypatch.setenv("MM2B200_FILL_CTAS", str(ctas))
                monkeypatch.setenv("MM2B200_RING", str(ring))
                with api.ChainContext(device=0, n_slots=slots, slot_anchors=slot_anchors) as c:
                    for _ in range(2):
                        assert _digest(c.chain_batch(prm, offs, a)) == want, (ctas, ring, slots)
                        runs += 1
    assert runs >= 50

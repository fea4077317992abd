"""world_size-2 (and 3) runs over gloo on the CPU: the N>1 host logic -- shard selection with the C partitioner,
per-rank work on the rank's own reads, result gathering in submission order.  The per-rank chaining itself needs a GPU,
so here the oracle stands in for it (tests are allowed to); the GPU version of the same flow is test_multi_gpu.py."""
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_reads, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle
    from minimap2_chaindp_b200 import api, shard, workload
    spec = workload.CONFIGS["map-ont-100k"]
    offs = workload.count(spec, 0, n_reads)                       # every rank knows the read sizes, not the anchors
    r0, r1 = shard.shard_of(offs, rank, world)
    lo, la = workload.generate(spec, r0, r1 - r0)                 # a rank materialises only its own shard
    res = oracle.chain_batch(oracle.params(spec.preset), lo, la)
    local = api.BatchResult(res.n_u, res.u_off, res.b_off, res.u, res.b, n_visited=res.visited)
    merged = shard.gather_results(local, dst=0)
    if rank == 0:
        full_offs, full_a = workload.generate(spec, 0, n_reads)
        want = oracle.chain_batch(oracle.params(spec.preset), full_offs, full_a)
        ok = all(np.array_equal(getattr(merged, k), getattr(want, k)) for k in ("n_u", "u_off", "b_off", "u", "b"))
        q.put((ok, merged.n_visited == want.visited, int(merged.u_off[-1])))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_run_equals_single_run(world):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, 400, q)) for r in range(world)]
    [p.start() for p in procs]
    ok, vis_ok, n_chains = q.get(timeout=180)
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    assert ok and vis_ok and n_chains > 0

"""mm2b200_sketch_device_noalloc (include/mm2sketch_b200.h): the sketching kernels with their scratch in a static device array.
Collected last on purpose.  Through the C ABI, against the existing entry point mm2b200_sketch_device on the same device
buffers: bit-exact (integer work), on two calls in a row that share the scratch array (the second batch is smaller than the
first, with other parameters, so anything the first call left behind would show)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_READS, MAX_BASES = 4096, 1 << 20


def _batch(rng, n_reads, max_len):
    acgt = np.frombuffer(b"ACGTacgtN", np.uint8)
    lens = rng.integers(0, max_len, n_reads)
    lens[rng.integers(0, n_reads)] = 0                              # an empty read
    off = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    p = np.array([0.24, 0.24, 0.24, 0.24, 0.008, 0.008, 0.008, 0.008, 0.008])
    seq = acgt[rng.choice(9, int(off[-1]), p=p / p.sum())]
    return off, seq


def _both(ctx, off, seq, w, k, hpc, rid=None, mv_cap=None):
    import torch
    from minimap2_chaindp_b200 import api, sketch
    L = sketch._lib()
    dev = torch.device("cuda", 0)
    n_reads, n_bases = len(off) - 1, int(off[-1])
    d_off = torch.from_numpy(off).to(dev)
    d_seq = torch.from_numpy(seq).to(dev) if n_bases else torch.zeros(1, dtype=torch.uint8, device=dev)
    d_rid = torch.from_numpy(rid.astype(np.uint32).view(np.int32)).to(dev) if rid is not None else None
    prm = sketch.SketchParams(w, k, hpc)
    # existing entry point
    out, n_old = sketch._DevSketch(), C.c_int64()
    rc = L.mm2b200_sketch_device(ctx._h, C.byref(prm), n_reads, n_bases, d_off.data_ptr(), d_seq.data_ptr(), d_rid.data_ptr() if d_rid is not None else None,
                                 None, C.byref(out), C.byref(n_old))
    assert rc == 0, api.last_error()
    torch.cuda.synchronize()
    from cuda.bindings import runtime as rt
    want_off = np.zeros(n_reads + 1, np.int64)
    want = np.zeros((n_old.value, 2), np.uint64)
    assert rt.cudaMemcpy(want_off.ctypes.data, out.d_mv_off, want_off.nbytes, rt.cudaMemcpyKind.cudaMemcpyDeviceToHost)[0] == rt.cudaError_t.cudaSuccess
    if n_old.value:
        assert rt.cudaMemcpy(want.ctypes.data, out.d_mv, want.nbytes, rt.cudaMemcpyKind.cudaMemcpyDeviceToHost)[0] == rt.cudaError_t.cudaSuccess
    # the allocation-free entry point: every buffer is the caller's
    cap = n_bases if mv_cap is None else mv_cap
    d_mv = torch.full((max(cap, 1), 2), -1, dtype=torch.int64, device=dev)
    d_mv_off = torch.full((n_reads + 1,), -1, dtype=torch.int64, device=dev)
    d_tot = torch.full((8,), -1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    n_new = C.c_int64(-1)
    rc = L.mm2b200_sketch_device_noalloc(C.byref(prm), n_reads, n_bases, d_off.data_ptr(), d_seq.data_ptr(), d_rid.data_ptr() if d_rid is not None else None,
                                         None, d_mv.data_ptr(), cap, d_mv_off.data_ptr(), d_tot.data_ptr(), C.byref(n_new))
    return rc, n_new.value, d_mv, d_mv_off, d_tot, want_off, want


@pytest.fixture(scope="module")
def ctx():
    from minimap2_chaindp_b200 import api
    c = api.ChainContext(device=0)
    yield c
    c.close()


def test_two_calls_in_a_row_match_the_allocating_entry_point(ctx):
    from minimap2_chaindp_b200 import api
    rng = np.random.default_rng(2024)
    calls = [(_batch(rng, 300, 6000), 10, 15, 0, True),              # k_sketch_odd<10>, ~0.9 M bases: most of the scratch
             (_batch(rng, 40, 3000), 5, 19, 1, False),               # general kernel with HPC, a small batch over the leftovers
             (_batch(rng, 64, 2500), 7, 13, 0, False)]               # k_sketch_odd<0>
    for (off, seq), w, k, hpc, with_rid in calls:
        assert off[-1] <= MAX_BASES
        rid = (np.arange(len(off) - 1) % 11) if with_rid else None
        rc, n_new, d_mv, d_mv_off, d_tot, want_off, want = _both(ctx, off, seq, w, k, hpc, rid)
        assert rc == 0, api.last_error()
        assert n_new == want.shape[0] and n_new > 0
        tot = d_tot.cpu().numpy()
        assert tot[0] == n_new and tot[1] == 0
        assert np.array_equal(d_mv_off.cpu().numpy(), want_off)
        got = d_mv[:n_new].cpu().numpy().view(np.uint64)
        assert np.array_equal(got, want), f"w={w} k={k} hpc={hpc}: minimizers differ from mm2b200_sketch_device"


def test_limits_are_rejected_and_nothing_is_allocated(ctx):
    import torch
    from minimap2_chaindp_b200 import api, sketch
    L = sketch._lib()
    prm = sketch.SketchParams(10, 15, 0)
    dev = torch.device("cuda", 0)
    d = torch.zeros(16, dtype=torch.int64, device=dev)
    n = C.c_int64()
    for n_reads, n_bases in ((MAX_READS + 1, 100), (10, MAX_BASES + 1)):
        rc = L.mm2b200_sketch_device_noalloc(C.byref(prm), n_reads, n_bases, d.data_ptr(), d.data_ptr(), None, None, d.data_ptr(), 1, d.data_ptr(), d.data_ptr(), C.byref(n))
        assert rc == api.ERR_ARG and "exceeds the static scratch" in api.last_error()
    # an output that is too small is reported, not overrun
    rng = np.random.default_rng(7)
    off, seq = _batch(rng, 20, 2000)
    rc, n_new, d_mv, d_mv_off, d_tot, want_off, want = _both(ctx, off, seq, 10, 15, 0, mv_cap=8)
    assert rc == api.ERR_ARG and n_new == want.shape[0] and int(d_tot.cpu()[1]) == 1
    assert bool((d_mv == -1).all())
    # device memory in use does not move across calls (torch's caching allocator hands the same blocks back)
    off, seq = _batch(rng, 50, 4000)
    _both(ctx, off, seq, 10, 15, 0)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(0)[0]
    for _ in range(3):
        rc = _both(ctx, off, seq, 10, 15, 0)[0]
        assert rc == 0
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(0)[0] == free0

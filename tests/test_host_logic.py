"""CPU-only checks of the boundary: the C-ABI library loads without a GPU and exports every symbol the header
declares; compute entry points fail loudly (no CPU fallback); the read partitioner; the product never touches oracle/."""
import os
import re

import numpy as np
import pytest

from minimap2_chaindp_b200 import api, shard, workload

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols(header="mm2chain_b200.h"):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mm2b200_[a-z0-9_]+|mm_chain_dp[a-z_]*)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    names = _declared_symbols()
    assert len(names) >= 18 and "mm_chain_dp" in names and "mm2b200_chain_batch" in names
    L = api.lib()
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/mm2chain_b200.h but not exported"
    assert sorted(api.EXPORTED_SYMBOLS) == names
    assert b"sm_100a" in L.mm2b200_version()


def test_library_exports_every_seeding_symbol():
    from minimap2_chaindp_b200 import seed
    names = _declared_symbols("mm2seed_b200.h")
    assert len(names) >= 10 and "mm2b200_index_load" in names and "mm2b200_seed_work" in names
    L = api.lib()
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/mm2seed_b200.h but not exported"
    assert sorted(seed.EXPORTED_SYMBOLS) == names
    import ctypes as C
    assert C.sizeof(seed.SeedParams) == 44


def test_library_exports_every_post_processing_symbol():
    from minimap2_chaindp_b200 import post
    import ctypes as C
    names = _declared_symbols("mm2post_b200.h")
    L = api.lib()
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/mm2post_b200.h but not exported"
    assert sorted(post.EXPORTED_SYMBOLS) == names
    assert C.sizeof(post.PostParams) == 36 and post.REG_DTYPE.itemsize == 80


def test_library_exports_every_sketching_symbol():
    from minimap2_chaindp_b200 import sketch
    names = _declared_symbols("mm2sketch_b200.h")
    L = api.lib()
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/mm2sketch_b200.h but not exported"
    assert sorted(sketch.EXPORTED_SYMBOLS) == names


def test_library_exports_the_composed_entry():
    from minimap2_chaindp_b200 import seed
    names = _declared_symbols("mm2map_b200.h")
    assert names == sorted(seed.MAP_EXPORTED_SYMBOLS)
    for n in names:
        assert hasattr(api.lib(), n)


def test_library_exports_every_driver_symbol():
    """include/mm2fpgadrv_b200.h: the 11 symbols the fork's objects expect from libfpgadrv.a (fpga.h:37-62)"""
    from minimap2_chaindp_b200 import fpgadrv
    src = open(os.path.join(ROOT, "include", "mm2fpgadrv_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = sorted(set(re.findall(r"\b(fpga_[a-z_]+|mm2b200_[a-z0-9_]+)\s*\(", src)))
    assert len([n for n in names if n.startswith("fpga_")]) == 11
    L = api.lib()
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/mm2fpgadrv_b200.h but not exported"
    assert sorted(fpgadrv.EXPORTED_SYMBOLS) == names
    # without fpga_init nothing is served: no buffer, no packet, no counters
    F = fpgadrv._lib()
    import ctypes as C
    n = C.c_int(5)
    assert not F.fpga_get_writebuf(64, 0) and not F.fpga_get_retbuf(C.byref(n), 3) and n.value == 0
    assert F.mm2b200_shim_stats(None, None, None, None, None) == -1


def test_index_streams_are_validated_without_a_gpu():
    """size checks of mm2b200_index_commit come before any device work; without a context nothing can be created"""
    import torch
    from minimap2_chaindp_b200 import seed
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    L = seed._lib()
    import ctypes as C
    h = C.c_void_p()
    assert L.mm2b200_index_create(None, C.byref(h)) == api.ERR_ARG
    assert L.mm2b200_index_load(None, None, 0, 4) == api.ERR_ARG


def test_struct_layouts_match_the_reference():
    assert api.SEED_DTYPE.itemsize == 24                      # struct new_seed, minimap.h:51-55
    assert [api.SEED_DTYPE.fields[k][1] for k in ("x", "y", "p", "f")] == [0, 8, 16, 20]
    import ctypes as C
    assert C.sizeof(api.Params) == 36


def test_no_cpu_fallback_without_gpu():
    """Without a CUDA device the context cannot be created and the drop-in reports an error instead of computing."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    assert api.lib().mm2b200_device_count() == 0
    with pytest.raises(api.ChainError) as ei:
        api.ChainContext(device=0)
    assert ei.value.code == api.ERR_CUDA
    a = np.array([[100, (15 << 32) | 10], [120, (15 << 32) | 30], [140, (15 << 32) | 50]], np.uint64)
    with pytest.raises(api.ChainError):
        api.mm_chain_dp(api.params("map-ont", min_sc=0, min_cnt=1), a)


def test_product_does_not_use_the_oracle():
    pkg = os.path.join(ROOT, "minimap2_chaindp_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".h", ".c", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text and "chain_oracle" not in text \
                    and "libmm2ref" not in text, f"{f} references the oracle"


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_partition_reads_balances_anchors(world):
    spec = workload.CONFIGS["map-ont-100k"]
    offs = workload.count(spec, 0, 5000)
    cuts = api.partition_reads(offs, world)
    assert cuts[0] == 0 and cuts[-1] == 5000 and np.all(np.diff(cuts) >= 0)
    load = np.array([offs[cuts[k + 1]] - offs[cuts[k]] for k in range(world)], np.float64)
    assert load.sum() == offs[-1]
    assert load.max() <= load.mean() * 1.02 + 2000           # within one read of perfect balance
    # absolute offsets (a shard of a larger array) and degenerate inputs
    cuts2 = api.partition_reads(offs[100:1101], world)
    assert cuts2[0] == 0 and cuts2[-1] == 1000
    assert api.partition_reads(np.zeros(1, np.int64), world).tolist() == [0] * (world + 1)
    assert api.partition_reads(np.zeros(11, np.int64), 2).tolist() == [0, 5, 10]


def test_merge_results_rebases_offsets():
    import oracle
    spec = workload.CONFIGS["map-ont-100k"]
    offs, a = workload.generate(spec, 0, 300)
    prm = oracle.params(spec.preset)
    whole = oracle.chain_batch(prm, offs, a)
    cuts = api.partition_reads(offs, 3)
    parts = []
    for k in range(3):
        o = offs[cuts[k]:cuts[k + 1] + 1]
        r = oracle.chain_batch(prm, o - o[0], a[o[0]:o[-1]])
        parts.append(api.BatchResult(r.n_u, r.u_off, r.b_off, r.u, r.b))
    merged = shard.merge_results(parts)
    for name in ("n_u", "u_off", "b_off", "u", "b"):
        assert np.array_equal(getattr(merged, name), getattr(whole, name)), name

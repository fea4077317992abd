"""minimap2_chaindp_b200 -- minimap2's anchor-chaining DP (chain.c of stormalex/minimap2_chaindp) on NVIDIA B200.

The product is ``libmm2chain_b200.so`` (CUDA kernels for sm_100a behind the C ABI in ``include/mm2chain_b200.h``);
this package holds its sources (``csrc/``), the build recipe (``build.py``), the ctypes mirror of the reference
interface (``api.py``), the synthetic workload generator (``workload.py``) and the read-batch sharding helpers
(``shard.py``).  There is no CPU implementation of the chaining path in here.
"""
from .api import (BatchResult, ChainContext, ChainError, Params, PinnedArray, PRESETS, SEED_DTYPE,  # noqa: F401
                  mm_chain_dp, mm_chain_dp_bottom, mm_chain_dp_fpga, params)

__all__ = ["BatchResult", "ChainContext", "ChainError", "Params", "PinnedArray", "PRESETS", "SEED_DTYPE",
           "mm_chain_dp", "mm_chain_dp_bottom", "mm_chain_dp_fpga", "params"]

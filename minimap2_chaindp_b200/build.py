"""Builds the native pieces in-tree (so the .so files travel to the GPU box with the snapshot).

* ``minimap2_chaindp_b200/libmm2chain_b200.so`` -- CUDA kernels + C ABI (nvcc, sm_100a only)
* ``minimap2_chaindp_b200/libmm2workload.so``   -- synthetic anchor generator (gcc, host only)
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libmm2chain_b200.so")
WL_LIB = os.path.join(PKG, "libmm2workload.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "--shared", "-Xcompiler", "-fPIC,-Wall,-Wno-unused-function", "-Xlinker", "-Bsymbolic-functions",
    "-cudart", "static",
]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: the CUDA extension cannot be built (there is no CPU fallback)")


def _stale(target: str, sources) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


CU_SOURCES = ("chain_kernels.cu", "chain_api.cu", "seed_kernels.cu", "seed_api.cu", "fpgadrv_shim.cu", "post.cu", "sketch.cu")
CU_HEADERS = ("chain_internal.h", "seed_internal.h", "sort_exact.cuh", "packet_format.h")
API_HEADERS = ("mm2chain_b200.h", "mm2seed_b200.h", "mm2fpgadrv_b200.h", "mm2post_b200.h", "mm2sketch_b200.h", "mm2map_b200.h")
OBJ_DIR = os.path.join(PKG, "_obj")


def build_cuda(force: bool = False, verbose: bool = False, extra=()) -> str:
    """Every .cu is compiled to its own object (in parallel, only when it or a header changed), then linked."""
    from concurrent.futures import ThreadPoolExecutor
    srcs = [os.path.join(CSRC, f) for f in CU_SOURCES]
    hdrs = [os.path.join(CSRC, f) for f in CU_HEADERS] + [os.path.join(ROOT, "include", f) for f in API_HEADERS]
    os.makedirs(OBJ_DIR, exist_ok=True)
    tag = os.path.join(OBJ_DIR, "flags.txt")
    flags = " ".join([*NVCC_FLAGS, *extra])
    if not os.path.exists(tag) or open(tag).read() != flags:
        force = True
    compile_flags = [f for f in NVCC_FLAGS if f not in ("--shared",)]
    for drop in ("-Xlinker", "-Bsymbolic-functions", "-cudart", "static"):
        compile_flags.remove(drop)
    jobs = []
    for src in srcs:
        obj = os.path.join(OBJ_DIR, os.path.basename(src)[:-3] + ".o")
        if force or _stale(obj, [src, *hdrs]):
            cmd = [_nvcc(), *compile_flags, *extra, "-I", os.path.join(ROOT, "include"), "-I", CSRC, "-c", "-o", obj, src]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
                print(" ".join(cmd), file=sys.stderr)
            jobs.append(cmd)
    if jobs:
        with ThreadPoolExecutor(len(jobs)) as ex:
            for r in ex.map(lambda c: subprocess.run(c), jobs):
                if r.returncode:
                    raise subprocess.CalledProcessError(r.returncode, r.args)
    objs = [os.path.join(OBJ_DIR, f[:-3] + ".o") for f in CU_SOURCES]
    if jobs or force or _stale(LIB, objs):
        subprocess.run([_nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", "--shared", "-Xlinker", "-Bsymbolic-functions",
                        "-cudart", "static", "-o", LIB, *objs, "-lpthread"], check=True)
        open(tag, "w").write(flags)
    return LIB


def build_workload(force: bool = False) -> str:
    srcs = [os.path.join(CSRC, "workload.c"), os.path.join(CSRC, "workload_seed.c")]
    if force or _stale(WL_LIB, srcs):
        subprocess.run(["gcc", "-O2", "-fPIC", "-shared", "-fopenmp", "-Wall", "-o", WL_LIB, *srcs, "-lm", "-lpthread"], check=True)
    return WL_LIB


def build_all(force: bool = False, verbose: bool = False) -> None:
    build_cuda(force, verbose)
    build_workload(force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)

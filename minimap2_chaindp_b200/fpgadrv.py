"""ctypes stub over the driver boundary (include/mm2fpgadrv_b200.h): the fork's ``fpga_*`` symbols (fpga.h:37-62) as
``libmm2chain_b200.so`` exports them.  Holds no logic of its own; tests and ``bench.py`` use it the way the fork's
``map.c`` / ``fpga_chaindp.c`` use ``libfpgadrv.a``:

* ``fpga_init`` / ``fpga_finalize``                        main.c:512,614
* ``fpga_load_index`` (chunked) / ``fpga_set_params``      main.c:201-204,243
* ``fpga_get_writebuf_thread`` -> copy -> ``fpga_writebuf_submit``   map.c:439-444
* ``fpga_get_retbuf`` -> copy -> ``fpga_release_retbuf``   fpga_chaindp.c:241-262
* ``fpga_exit_block`` / ``fpga_set_block``                 main.c:608,613
"""
from __future__ import annotations

import ctypes as C
import time

import numpy as np

from . import api

TYPE_CD, RET_TYPE_CS, BUF_TYPE_SW = 1, 3, 0
TYPE_INDEX_B, TYPE_INDEX_H, TYPE_INDEX_V, TYPE_INDEX_P = 4, 5, 6, 7

# every symbol include/mm2fpgadrv_b200.h declares
EXPORTED_SYMBOLS = (
    "fpga_init", "fpga_finalize", "fpga_set_params", "fpga_load_index", "fpga_get_writebuf", "fpga_get_writebuf_thread",
    "fpga_writebuf_submit", "fpga_get_retbuf", "fpga_release_retbuf", "fpga_exit_block", "fpga_set_block",
    "mm2b200_shim_stats",
)

_bound = False


def _lib():
    global _bound
    L = api.lib()
    if not _bound:
        vp, i64 = C.c_void_p, C.c_int64
        L.fpga_init.argtypes = [C.c_int]
        L.fpga_finalize.restype = None
        L.fpga_set_params.argtypes = [C.c_int] * 6
        L.fpga_set_params.restype = None
        L.fpga_load_index.argtypes = [vp, C.c_int, C.c_int]
        L.fpga_load_index.restype = None
        L.fpga_get_writebuf.argtypes = [C.c_ulong, C.c_int]
        L.fpga_get_writebuf.restype = vp
        L.fpga_get_writebuf_thread.argtypes = [C.c_ulong, C.c_int, C.c_int]
        L.fpga_get_writebuf_thread.restype = vp
        L.fpga_writebuf_submit.argtypes = [vp, C.c_uint, C.c_uint]
        L.fpga_get_retbuf.argtypes = [C.POINTER(C.c_int), C.c_int]
        L.fpga_get_retbuf.restype = vp
        L.fpga_release_retbuf.argtypes = [vp]
        L.fpga_exit_block.restype = None
        L.fpga_set_block.restype = None
        L.mm2b200_shim_stats.argtypes = [C.POINTER(i64)] * 5
        _bound = True
    return L


def init(flag: int = 0) -> None:
    if _lib().fpga_init(flag) != 0:
        raise api.ChainError(api.ERR_CUDA, api.last_error() or "fpga_init failed")


def finalize() -> None:
    _lib().fpga_finalize()


def set_params(bw, is_cdna, max_skip, min_sc, flag, max_occ) -> None:
    _lib().fpga_set_params(int(bw), int(is_cdna), int(max_skip), int(min_sc), int(flag), int(max_occ))


def load_index(stream, type_: int, chunk: int = 1 << 30) -> None:
    """One index stream in chunks, like load_index_buf (main.c:179)."""
    raw = np.ascontiguousarray(stream).view(np.uint8).reshape(-1)
    for at in range(0, max(raw.size, 1), chunk):
        part = raw[at:at + chunk]
        if part.size:
            _lib().fpga_load_index(part.ctypes.data, int(part.size), type_)


def submit(packet: bytes, tid: int = 0) -> None:
    """map.c:438-444: wait for a write buffer, copy the packet in, hand it over."""
    L = _lib()
    n = len(packet)
    while True:
        buf = L.fpga_get_writebuf_thread(n, BUF_TYPE_SW, tid)
        if buf:
            break
        time.sleep(50e-6)
    C.memmove(buf, packet, n)
    if L.fpga_writebuf_submit(buf, n, TYPE_CD) != 0:
        raise api.ChainError(api.ERR_STATE, "fpga_writebuf_submit failed")


def receive():
    """fpga_chaindp.c:241-262: the next response packet as bytes, or None once fpga_exit_block() has released the wait and
    nothing is left."""
    L = _lib()
    n = C.c_int(0)
    buf = L.fpga_get_retbuf(C.byref(n), RET_TYPE_CS)
    if n.value == 0:
        return None
    out = C.string_at(buf, n.value)
    L.fpga_release_retbuf(buf)
    return out


def exit_block() -> None:
    _lib().fpga_exit_block()


def set_block() -> None:
    _lib().fpga_set_block()


def stats() -> dict:
    v = [C.c_int64() for _ in range(5)]
    if _lib().mm2b200_shim_stats(*[C.byref(x) for x in v]) != 0:
        raise api.ChainError(api.ERR_STATE, "fpga_init has not run")
    return dict(zip(("packets", "reads", "batches", "err_reads", "launches"), (x.value for x in v)))

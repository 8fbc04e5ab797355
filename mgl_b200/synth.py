"""Synthetic benchmark pairs (mgl_b200/csrc/sw_synth.cpp, include/mgl_synth.h) as numpy arrays.  The generator
is its own small library, libmgl_synth.so: importing this module does not load libmgl_sw.so (only pinned=True
does, for cudaHostAlloc), so the reference arm of bench.py maps nothing of the product."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libmgl_synth.so")
_lib = None


def use_library(path: str):
    """Load the generator from another build of the same source (bench.py's reference arm uses oracle/'s copy so that
    it maps nothing of the product tree).  Must be called before the first generate()."""
    global LIB_PATH, _lib
    LIB_PATH, _lib = path, None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} not built (run __graft_entry__.build() or make -C mgl_b200/csrc)")
        L = C.CDLL(LIB_PATH)
        L.mgl_sw_synth_sizes.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.c_int64, C.c_void_p, C.c_void_p]
        L.mgl_sw_synth_fill.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib

SEEDS = {2: 0xC2, 3: 0xC3, 4: 0xC4, 5: 0xC5}
SCORES = {2: (10, -15, -30, -5), 3: (200, -150, -260, -11), 4: (200, -150, -260, -11), 5: (10, -15, -30, -5)}


def generate(config: int, n: int, first: int = 0, seed: int | None = None, pinned: bool = False, threads: int = 1):
    """-> dict(arena, ref_off, ref_len, alt_off, alt_len, strategies, params, cells)"""
    L = lib()
    seed = SEEDS[config] if seed is None else seed
    ref_len = np.zeros(n, dtype=np.uint32)
    alt_len = np.zeros(n, dtype=np.uint32)
    # a pair is a pure function of its index, so ranges of the batch are generated side by side (ctypes drops the GIL)
    T = max(1, min(int(threads), n // 65536 + 1))
    cuts = [n * k // T for k in range(T + 1)]

    def in_ranges(fn):
        import threading
        rcs = [0] * T
        def work(k):
            rcs[k] = fn(cuts[k], cuts[k + 1] - cuts[k])
        th = [threading.Thread(target=work, args=(k,)) for k in range(1, T)]
        for t in th:
            t.start()
        work(0)
        for t in th:
            t.join()
        return max(abs(r) for r in rcs)

    rc = in_ranges(lambda lo, cnt: L.mgl_sw_synth_sizes(config, seed, first + lo, cnt, ref_len.ctypes.data + 4 * lo, alt_len.ctypes.data + 4 * lo))
    if rc:
        raise ValueError(f"mgl_sw_synth_sizes({config}, n={n}) failed")
    sizes = np.empty(2 * n, dtype=np.uint64)
    sizes[0::2], sizes[1::2] = ref_len, alt_len
    offs = np.zeros(2 * n, dtype=np.uint64)
    offs[1:] = np.cumsum(sizes)[:-1]
    total = int(sizes.sum())
    if pinned:
        from . import native
        arena = native.pinned_array(total, np.uint8)
    else:
        arena = np.zeros(total, dtype=np.uint8)
    ref_off, alt_off = offs[0::2].copy(), offs[1::2].copy()
    rc = in_ranges(lambda lo, cnt: L.mgl_sw_synth_fill(config, seed, first + lo, cnt, ref_off.ctypes.data + 8 * lo, alt_off.ctypes.data + 8 * lo, arena.ctypes.data))
    if rc:
        raise ValueError(f"mgl_sw_synth_fill({config}, n={n}) failed")
    idx = np.arange(first, first + n)
    if config == 3:
        strategies = np.where(idx % 2 == 0, 1, 2).astype(np.int32)     # even SOFTCLIP, odd INDEL
    else:
        strategies = np.full(n, 1, dtype=np.int32)
    cells = int((ref_len.astype(np.uint64) * alt_len.astype(np.uint64)).sum())
    return dict(arena=arena, ref_off=ref_off, ref_len=ref_len, alt_off=alt_off, alt_len=alt_len,
                strategies=strategies, params=SCORES[config], cells=cells)


def pair_bytes(d, p):
    a = d["arena"]
    r = bytes(a[int(d["ref_off"][p]):int(d["ref_off"][p]) + int(d["ref_len"][p])])
    q = bytes(a[int(d["alt_off"][p]):int(d["alt_off"][p]) + int(d["alt_len"][p])])
    return r, q

"""Synthetic benchmark pairs (mgl_b200/csrc/sw_synth.cpp) as numpy arrays."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import native

SEEDS = {2: 0xC2, 3: 0xC3, 4: 0xC4, 5: 0xC5}
SCORES = {2: (10, -15, -30, -5), 3: (200, -150, -260, -11), 4: (200, -150, -260, -11), 5: (10, -15, -30, -5)}


def generate(config: int, n: int, first: int = 0, seed: int | None = None, pinned: bool = False):
    """-> dict(arena, ref_off, ref_len, alt_off, alt_len, strategies, params, cells)"""
    L = native.lib()
    L.mgl_sw_synth_sizes.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.c_int64, C.c_void_p, C.c_void_p]
    L.mgl_sw_synth_fill.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
    seed = SEEDS[config] if seed is None else seed
    ref_len = np.zeros(n, dtype=np.uint32)
    alt_len = np.zeros(n, dtype=np.uint32)
    rc = L.mgl_sw_synth_sizes(config, seed, first, n, ref_len.ctypes.data, alt_len.ctypes.data)
    if rc:
        raise native.MglSwError(rc)
    sizes = np.empty(2 * n, dtype=np.uint64)
    sizes[0::2], sizes[1::2] = ref_len, alt_len
    offs = np.zeros(2 * n, dtype=np.uint64)
    offs[1:] = np.cumsum(sizes)[:-1]
    total = int(sizes.sum())
    arena = native.pinned_array(total, np.uint8) if pinned else np.zeros(total, dtype=np.uint8)
    ref_off, alt_off = offs[0::2].copy(), offs[1::2].copy()
    rc = L.mgl_sw_synth_fill(config, seed, first, n, ref_off.ctypes.data, alt_off.ctypes.data, arena.ctypes.data)
    if rc:
        raise native.MglSwError(rc)
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

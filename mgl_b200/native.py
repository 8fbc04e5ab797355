"""ctypes binding of libmgl_sw.so (include/mgl_sw.h).  No CPU fallback: if the
library is missing or no B200 is usable, calls raise."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# MGL_SW_LIB: developer knob for A/B runs of differently built kernels; unset in production
LIB_PATH = os.environ.get("MGL_SW_LIB") or os.path.join(HERE, "libmgl_sw.so")

OK, E_INVALID, E_SCORES, E_PAIR, E_CUDA, E_OPS_CAPACITY, E_NO_DEVICE = 0, -1, -2, -3, -4, -5, -6
_CODES = {E_INVALID: "invalid argument", E_SCORES: "score set out of range", E_PAIR: "empty or over-long sequence",
          E_CUDA: "CUDA error", E_OPS_CAPACITY: "ops buffer too small", E_NO_DEVICE: "no usable B200 (sm_100a) device"}


class MglSwError(RuntimeError):
    def __init__(self, code, detail=""):
        self.code = code
        super().__init__(f"mgl_sw error {code} ({_CODES.get(code, '?')}){': ' + detail if detail else ''}")


class Batch(C.Structure):
    _fields_ = [("n", C.c_int64), ("arena", C.c_void_p), ("arena_bytes", C.c_uint64),
                ("ref_off", C.c_void_p), ("ref_len", C.c_void_p), ("alt_off", C.c_void_p), ("alt_len", C.c_void_p),
                ("strategies", C.c_void_p), ("strategy", C.c_int32),
                ("match", C.c_int32), ("mismatch", C.c_int32), ("gap_open", C.c_int32), ("gap_extend", C.c_int32)]


class Result(C.Structure):
    _fields_ = [("offsets", C.c_void_p), ("cigar_start", C.c_void_p), ("ops", C.c_void_p),
                ("ops_cap", C.c_uint64), ("ops_total", C.c_uint64)]


class TextResult(C.Structure):
    _fields_ = [("offsets", C.c_void_p), ("text_start", C.c_void_p), ("text", C.c_void_p),
                ("text_cap", C.c_uint64), ("text_total", C.c_uint64)]


class Stats(C.Structure):
    _fields_ = [(k, C.c_uint64) for k in ("cells", "pairs", "pairs_s16", "pairs_s32", "fill_launches", "other_launches",
                                          "chunks", "trace_bytes", "h2d_bytes", "d2h_bytes")] + \
               [(k, C.c_double) for k in ("fill_ms", "trace_ms", "post_ms", "device_ms")]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MglSwError(E_NO_DEVICE, f"{LIB_PATH} not built (run __graft_entry__.build() or make -C mgl_b200/csrc)")
        L = C.CDLL(LIB_PATH)
        L.mgl_sw_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
        L.mgl_sw_create_multi.argtypes = [C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_void_p)]
        L.mgl_sw_handle_devices.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.c_int]
        L.mgl_sw_default_handle.argtypes = [C.POINTER(C.c_void_p)]
        L.mgl_sw_default_release.restype = None
        L.mgl_sw_shutdown.restype = None
        L.mgl_sw_destroy.argtypes = [C.c_void_p]
        L.mgl_sw_destroy.restype = None
        L.mgl_sw_last_error.argtypes = [C.c_void_p]
        L.mgl_sw_last_error.restype = C.c_char_p
        L.mgl_sw_version.restype = C.c_char_p
        L.mgl_sw_set_option.argtypes = [C.c_void_p, C.c_char_p, C.c_int64]
        L.mgl_sw_align_batch.argtypes = [C.c_void_p, C.POINTER(Batch), C.POINTER(Result), C.POINTER(Stats)]
        L.mgl_sw_align_batch_text.argtypes = [C.c_void_p, C.POINTER(Batch), C.POINTER(TextResult), C.POINTER(Stats)]
        L.mgl_sw_align_batch_text.restype = C.c_int
        L.mgl_sw_score_batch.argtypes = [C.c_void_p, C.POINTER(Batch), C.c_void_p, C.POINTER(Stats)]
        L.mgl_sw_score_batch.restype = C.c_int
        L.mgl_sw_resident_prepare.argtypes = [C.c_void_p, C.POINTER(Batch)]
        L.mgl_sw_resident_run.argtypes = [C.c_void_p, C.POINTER(Stats)]
        L.mgl_sw_resident_fetch.argtypes = [C.c_void_p, C.POINTER(Result)]
        L.mgl_sw_resident_drop.argtypes = [C.c_void_p]
        L.mgl_sw_resident_drop.restype = None
        L.mgl_sw_align_one.argtypes = [C.c_char_p, C.c_int32, C.c_char_p, C.c_int32] + [C.c_int32] * 5 + \
                                      [C.c_char_p, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
        L.mgl_sw_format_cigar.argtypes = [C.c_void_p, C.c_int64, C.c_char_p, C.c_int64]
        L.mgl_sw_host_alloc.argtypes = [C.c_uint64]
        L.mgl_sw_host_alloc.restype = C.c_void_p
        L.mgl_sw_host_free.argtypes = [C.c_void_p]
        L.mgl_sw_host_free.restype = None
        L.mgl_sw_pack2bit_host.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_int]
        L.mgl_sw_plan_summary.argtypes = [C.POINTER(Batch), C.c_int, C.c_void_p]
        L.mgl_sw_int_peak_name.argtypes = [C.c_int]
        L.mgl_sw_int_peak_name.restype = C.c_char_p
        L.mgl_sw_int_peak.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        _lib = L
    return _lib


def pinned_array(shape, dtype):
    """numpy array over cudaHostAlloc'd memory (kept alive by the array's base object)."""
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) if not isinstance(shape, int) else int(shape)
    nbytes = max(n * dtype.itemsize, 1)
    ptr = lib().mgl_sw_host_alloc(nbytes)
    if not ptr:
        raise MglSwError(E_CUDA, "cudaHostAlloc failed")

    class _Owner:
        def __init__(self, p):
            self.p = p

        def __del__(self):
            try:
                lib().mgl_sw_host_free(self.p)
            except Exception:
                pass

    buf = (C.c_uint8 * nbytes).from_address(ptr)
    buf._owner = _Owner(ptr)
    return np.frombuffer(buf, dtype=dtype, count=n).reshape(shape)


def make_batch(arena, ref_off, ref_len, alt_off, alt_len, params, strategy=1, strategies=None):
    """Build a Batch over numpy arrays; returns (Batch, keepalive tuple)."""
    arena = np.ascontiguousarray(arena, dtype=np.uint8)
    ref_off = np.ascontiguousarray(ref_off, dtype=np.uint64)
    alt_off = np.ascontiguousarray(alt_off, dtype=np.uint64)
    ref_len = np.ascontiguousarray(ref_len, dtype=np.uint32)
    alt_len = np.ascontiguousarray(alt_len, dtype=np.uint32)
    st = None
    if strategies is not None:
        st = np.ascontiguousarray(strategies, dtype=np.int32)
    b = Batch()
    b.n = len(ref_len)
    b.arena = arena.ctypes.data
    b.arena_bytes = arena.nbytes
    b.ref_off, b.ref_len = ref_off.ctypes.data, ref_len.ctypes.data
    b.alt_off, b.alt_len = alt_off.ctypes.data, alt_len.ctypes.data
    b.strategies = st.ctypes.data if st is not None else None
    b.strategy = int(strategy)
    b.match, b.mismatch, b.gap_open, b.gap_extend = (int(x) for x in params)
    return b, (arena, ref_off, ref_len, alt_off, alt_len, st)

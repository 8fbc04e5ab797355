"""Host-side mirror of the reference's Java binding
(src/main/java/com/microsoft/mgl/smithwaterman/MicrosoftSmithWaterman.java:20-97),
same names and argument meaning, over the C ABI of libmgl_sw.so -- plus the
batched call the GPU needs.  This image has no JDK, so Python stands where the
Java class would; the JNI symbols themselves are in the same library."""
from __future__ import annotations

import ctypes as C
import enum
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import native

OP_CHARS = "MIDNSHP=X"


class SWOverhangStrategy(enum.IntEnum):
    """gatk-native-bindings SWOverhangStrategy; values as getStrategy() maps them (.java:39-56)."""
    SOFTCLIP = 0x01
    INDEL = 0x02
    LEADING_INDEL = 0x04
    IGNORE = 0x08


@dataclass(frozen=True)
class SWParameters:
    """gatk-native-bindings SWParameters: match value, mismatch penalty, gap open / extend penalties."""
    match_value: int
    mismatch_penalty: int
    gap_open_penalty: int
    gap_extend_penalty: int

    def as_tuple(self):
        return (self.match_value, self.mismatch_penalty, self.gap_open_penalty, self.gap_extend_penalty)


@dataclass(frozen=True)
class SWNativeAlignerResult:
    cigar: str
    alignment_offset: int


@dataclass
class BatchResult:
    offsets: np.ndarray        # int32 [n]
    cigar_start: np.ndarray    # uint64 [n + 1]
    ops: np.ndarray            # uint32 [total], (len << 4) | op
    stats: dict

    def cigar(self, p: int) -> str:
        o = self.ops[int(self.cigar_start[p]):int(self.cigar_start[p + 1])]
        return "".join(f"{int(x) >> 4}{OP_CHARS[int(x) & 15]}" for x in o)

    def cigars(self) -> List[str]:
        return [self.cigar(p) for p in range(len(self.offsets))]


def _params(parameters) -> Tuple[int, int, int, int]:
    return parameters.as_tuple() if isinstance(parameters, SWParameters) else tuple(int(x) for x in parameters)


def _pack_pairs(pairs: Sequence[Tuple[bytes, bytes]]):
    """[(ref, alt), ...] -> one arena, ref ‖ alt per pair (the layout of alignNative's readsBuffer, .java:73-75)"""
    n = len(pairs)
    ref_len = np.fromiter((len(r) for r, _ in pairs), dtype=np.uint32, count=n)
    alt_len = np.fromiter((len(a) for _, a in pairs), dtype=np.uint32, count=n)
    sizes = np.empty(2 * n, dtype=np.uint64)
    sizes[0::2], sizes[1::2] = ref_len, alt_len
    offs = np.zeros(2 * n, dtype=np.uint64)
    if n:
        offs[1:] = np.cumsum(sizes)[:-1]
    arena = np.frombuffer(b"".join(r + a for r, a in pairs), dtype=np.uint8)
    return arena, offs[0::2].copy(), ref_len, offs[1::2].copy(), alt_len


class MicrosoftSmithWaterman:
    """SWAlignerNativeBinding implementation backed by one B200, or by several: `device` may be a list, and every
    batch call is then cut by DP cells across those GPUs (mgl_sw_create_multi) and gathered in input order."""

    def __init__(self, device=0):
        self._device = device
        self._h = None

    # -- lifecycle (load() mirrors MicrosoftSmithWaterman.load, .java:34-37: False on any failure) --
    def load(self, temp_dir=None) -> bool:
        try:
            self._ensure()
            return True
        except native.MglSwError:
            return False

    def _ensure(self):
        if self._h is None:
            h = C.c_void_p()
            if isinstance(self._device, (list, tuple)):
                devs = (C.c_int * len(self._device))(*[int(x) for x in self._device])
                rc = native.lib().mgl_sw_create_multi(devs, len(self._device), C.byref(h))
            else:
                rc = native.lib().mgl_sw_create(int(self._device), C.byref(h))
            if rc != native.OK:
                raise native.MglSwError(rc)
            self._h = h
        return self._h

    def close(self):
        if self._h is not None:
            native.lib().mgl_sw_destroy(self._h)
            self._h = None

    def __enter__(self):
        self._ensure()
        return self

    def __exit__(self, *a):
        self.close()

    def set_option(self, name: str, value: int):
        rc = native.lib().mgl_sw_set_option(self._ensure(), name.encode(), int(value))
        if rc != native.OK:
            raise native.MglSwError(rc, name)

    def _err(self, rc):
        return native.MglSwError(rc, native.lib().mgl_sw_last_error(self._h).decode(errors="replace"))

    # -- the reference's call (.java:66-86) --
    def align(self, ref_array: bytes, alt_array: bytes, parameters, overhang_strategy) -> SWNativeAlignerResult:
        res = self.align_batch([(bytes(ref_array), bytes(alt_array))], parameters, overhang_strategy)
        return SWNativeAlignerResult(res.cigar(0), int(res.offsets[0]))

    # -- the batched call --
    def align_batch(self, pairs: Sequence[Tuple[bytes, bytes]], parameters, overhang_strategy=SWOverhangStrategy.SOFTCLIP,
                    strategies: Optional[Sequence[int]] = None) -> BatchResult:
        arena, ref_off, ref_len, alt_off, alt_len = _pack_pairs(pairs)
        return self.align_arrays(arena, ref_off, ref_len, alt_off, alt_len, parameters, overhang_strategy, strategies)

    def align_arrays(self, arena, ref_off, ref_len, alt_off, alt_len, parameters,
                     overhang_strategy=SWOverhangStrategy.SOFTCLIP, strategies=None, ops_cap: Optional[int] = None,
                     out: Optional[Tuple[np.ndarray, np.ndarray, np.ndarray]] = None) -> BatchResult:
        """Arena form: what a JNI/FFI caller passes.  `out` = (offsets, cigar_start, ops) preallocated (e.g. pinned)."""
        h = self._ensure()
        b, keep = native.make_batch(arena, ref_off, ref_len, alt_off, alt_len, _params(parameters),
                                    int(overhang_strategy), strategies)
        n = int(b.n)
        if out is None:
            cap = int(ops_cap) if ops_cap is not None else max(16, 8 * n)
            out = (np.zeros(n, dtype=np.int32), np.zeros(n + 1, dtype=np.uint64), np.zeros(cap, dtype=np.uint32))
        offsets, cigar_start, ops = out
        st = native.Stats()
        while True:
            r = native.Result()
            r.offsets, r.cigar_start, r.ops = offsets.ctypes.data, cigar_start.ctypes.data, ops.ctypes.data
            r.ops_cap = len(ops)
            rc = native.lib().mgl_sw_align_batch(h, C.byref(b), C.byref(r), C.byref(st))
            if rc == native.E_OPS_CAPACITY:
                ops = np.zeros(int(r.ops_total), dtype=np.uint32)
                continue
            if rc != native.OK:
                raise self._err(rc)
            break
        del keep
        return BatchResult(offsets, cigar_start, ops[: int(r.ops_total)], st.as_dict())

    # -- text CIGARs formatted on the device --
    def align_arrays_text(self, arena, ref_off, ref_len, alt_off, alt_len, parameters,
                          overhang_strategy=SWOverhangStrategy.SOFTCLIP, strategies=None, text_cap: Optional[int] = None,
                          out: Optional[Tuple[np.ndarray, np.ndarray, np.ndarray]] = None):
        """(offsets, text_start, text bytes, stats): pair p's CIGAR is text[text_start[p]:text_start[p + 1]] -- the
        characters alignNative writes to cigarBuffer (sw.cpp:251-252)."""
        h = self._ensure()
        b, keep = native.make_batch(arena, ref_off, ref_len, alt_off, alt_len, _params(parameters),
                                    int(overhang_strategy), strategies)
        n = int(b.n)
        if out is None:
            cap = int(text_cap) if text_cap is not None else max(64, 24 * n)
            out = (np.zeros(n, dtype=np.int32), np.zeros(n + 1, dtype=np.uint64), np.zeros(cap, dtype=np.uint8))
        offsets, text_start, text = out
        st = native.Stats()
        while True:
            r = native.TextResult()
            r.offsets, r.text_start, r.text = offsets.ctypes.data, text_start.ctypes.data, text.ctypes.data
            r.text_cap = len(text)
            rc = native.lib().mgl_sw_align_batch_text(h, C.byref(b), C.byref(r), C.byref(st))
            if rc == native.E_OPS_CAPACITY:
                text = np.zeros(int(r.text_total), dtype=np.uint8)
                continue
            if rc != native.OK:
                raise self._err(rc)
            break
        del keep
        return offsets, text_start, text[: int(r.text_total)], st.as_dict()

    def align_batch_text(self, pairs: Sequence[Tuple[bytes, bytes]], parameters, overhang_strategy=SWOverhangStrategy.SOFTCLIP,
                         strategies: Optional[Sequence[int]] = None):
        """[(cigar text, offset), ...] in input order."""
        arena, ref_off, ref_len, alt_off, alt_len = _pack_pairs(pairs)
        offsets, start, text, _ = self.align_arrays_text(arena, ref_off, ref_len, alt_off, alt_len, parameters, overhang_strategy, strategies)
        raw = text.tobytes()
        return [(raw[int(start[p]): int(start[p + 1])].decode("ascii"), int(offsets[p])) for p in range(len(pairs))]

    # -- score only: no traceback --
    def score_arrays(self, arena, ref_off, ref_len, alt_off, alt_len, parameters,
                     overhang_strategy=SWOverhangStrategy.SOFTCLIP, strategies=None, out: Optional[np.ndarray] = None):
        """(score3, stats): score3[p] = (H of the traceback's start cell, its ref row, its alt column), sw.cpp:155-170."""
        h = self._ensure()
        b, keep = native.make_batch(arena, ref_off, ref_len, alt_off, alt_len, _params(parameters),
                                    int(overhang_strategy), strategies)
        score3 = out if out is not None else np.zeros((int(b.n), 3), dtype=np.int32)
        st = native.Stats()
        rc = native.lib().mgl_sw_score_batch(h, C.byref(b), score3.ctypes.data, C.byref(st))
        if rc != native.OK:
            raise self._err(rc)
        del keep
        return score3, st.as_dict()

    def score_batch(self, pairs: Sequence[Tuple[bytes, bytes]], parameters, overhang_strategy=SWOverhangStrategy.SOFTCLIP,
                    strategies: Optional[Sequence[int]] = None) -> np.ndarray:
        arena, ref_off, ref_len, alt_off, alt_len = _pack_pairs(pairs)
        return self.score_arrays(arena, ref_off, ref_len, alt_off, alt_len, parameters, overhang_strategy, strategies)[0]

    # -- device-resident timing path used by bench.py --
    def resident_prepare(self, arena, ref_off, ref_len, alt_off, alt_len, parameters, overhang_strategy=1, strategies=None):
        b, keep = native.make_batch(arena, ref_off, ref_len, alt_off, alt_len, _params(parameters),
                                    int(overhang_strategy), strategies)
        rc = native.lib().mgl_sw_resident_prepare(self._ensure(), C.byref(b))
        if rc != native.OK:
            raise self._err(rc)
        self._resident_n = int(b.n)

    def resident_run(self) -> dict:
        st = native.Stats()
        rc = native.lib().mgl_sw_resident_run(self._ensure(), C.byref(st))
        if rc != native.OK:
            raise self._err(rc)
        return st.as_dict()

    def resident_fetch(self, ops_cap: Optional[int] = None) -> BatchResult:
        n = self._resident_n
        offsets, cigar_start = np.zeros(n, dtype=np.int32), np.zeros(n + 1, dtype=np.uint64)
        ops = np.zeros(int(ops_cap) if ops_cap else max(16, 8 * n), dtype=np.uint32)
        while True:
            r = native.Result()
            r.offsets, r.cigar_start, r.ops = offsets.ctypes.data, cigar_start.ctypes.data, ops.ctypes.data
            r.ops_cap = len(ops)
            rc = native.lib().mgl_sw_resident_fetch(self._ensure(), C.byref(r))
            if rc == native.E_OPS_CAPACITY:
                ops = np.zeros(int(r.ops_total), dtype=np.uint32)
                continue
            if rc != native.OK:
                raise self._err(rc)
            return BatchResult(offsets, cigar_start, ops[: int(r.ops_total)], {})

    def resident_drop(self):
        native.lib().mgl_sw_resident_drop(self._ensure())


def align_one_text(ref: bytes, alt: bytes, parameters, overhang_strategy) -> SWNativeAlignerResult:
    """Through mgl_sw_align_one, the entry the JNI alignNative symbol uses."""
    p = _params(parameters)
    cap = 12 * (len(ref) + len(alt)) + 16
    buf = C.create_string_buffer(cap)
    ln, off = C.c_int32(), C.c_int32()
    rc = native.lib().mgl_sw_align_one(ref, len(ref), alt, len(alt), p[0], p[1], p[2], p[3], int(overhang_strategy),
                                       buf, cap, C.byref(ln), C.byref(off))
    if rc != native.OK:
        raise native.MglSwError(rc)
    return SWNativeAlignerResult(buf.raw[: ln.value].decode(), off.value)

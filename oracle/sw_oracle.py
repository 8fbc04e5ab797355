"""ctypes loaders for the oracle (oracle/sw_oracle.c) and, when it has been
built, the reference compiled in place (oracle/_ref/libmgl_sw_ref.so).

TEST INFRASTRUCTURE: see oracle/__init__.py for who may import this.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import List, Optional, Sequence, Tuple

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "libsw_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libmgl_sw_ref.so")
SYNTH_SO = os.path.join(HERE, "libsynth_inputs.so")      # the benchmark input generator, for bench.py's reference arm

SOFTCLIP, INDEL, LEADING_INDEL, IGNORE = 1, 2, 4, 8      # sw_common.h:22-25
STRATEGIES = (SOFTCLIP, INDEL, LEADING_INDEL, IGNORE)
OP_CHARS = "MIDNS"


class OracleResult(C.Structure):
    _fields_ = [("offset", C.c_int32), ("n_ops", C.c_int32), ("score", C.c_int32),
                ("start_i", C.c_int32), ("start_j", C.c_int32)]


def build(force: bool = False) -> None:
    """Compile the C restatement (and the in-place reference when its sources are present)."""
    if force or not os.path.exists(ORACLE_SO) or not os.path.exists(SYNTH_SO) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(os.path.join(HERE, "sw_oracle.c")):
        subprocess.check_call(["make", "-C", HERE, "oracle"], stdout=subprocess.DEVNULL)
    ref_src = os.environ.get("MGL_REFERENCE", "/root/reference")
    if os.path.isdir(os.path.join(ref_src, "src/main/native/mgl_sw")):
        subprocess.check_call(["make", "-C", HERE, "ref", f"REFERENCE={ref_src}"], stdout=subprocess.DEVNULL)


_oracle = None
_ref = None


def _load_oracle():
    global _oracle
    if _oracle is None:
        if not os.path.exists(ORACLE_SO):
            build()
        lib = C.CDLL(ORACLE_SO)
        lib.oracle_sw_align.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                        C.c_int, C.POINTER(C.c_uint32), C.c_int, C.POINTER(OracleResult)]
        lib.oracle_sw_align.restype = C.c_int
        lib.oracle_sw_format.argtypes = [C.POINTER(C.c_uint32), C.c_int, C.c_char_p, C.c_int]
        lib.oracle_sw_format.restype = C.c_int
        _oracle = lib
    return _oracle


def have_ref() -> bool:
    return os.path.exists(REF_SO)


def _load_ref():
    global _ref
    if _ref is None:
        lib = C.CDLL(REF_SO)
        lib.mgl_ref_align.argtypes = [C.c_int, C.c_char_p, C.c_int, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                      C.c_int, C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_int32)]
        lib.mgl_ref_align.restype = C.c_int
        lib.mgl_ref_align_jni.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                          C.c_char_p]
        lib.mgl_ref_align_jni.restype = C.c_int
        lib.mgl_ref_max_threads.restype = C.c_int
        if hasattr(lib, "mgl_ref_scoremax"):
            lib.mgl_ref_scoremax.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                             C.c_int, C.POINTER(C.c_int32)]
            lib.mgl_ref_scoremax.restype = C.c_int
        lib.mgl_ref_align_batch.argtypes = [C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                            C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        lib.mgl_ref_align_batch.restype = C.c_double
        if hasattr(lib, "mgl_ref_check_batch"):
            lib.mgl_ref_check_batch.argtypes = [C.c_int, C.c_int64] + [C.c_void_p] * 5 + [C.c_int] * 5 + [C.c_void_p, C.c_int] + \
                                               [C.c_void_p] * 4 + [C.c_int, C.POINTER(C.c_double)]
            lib.mgl_ref_check_batch.restype = C.c_int64
        _ref = lib
    return _ref


def ops_to_text(ops: Sequence[int]) -> str:
    return "".join(f"{int(o) >> 4}{OP_CHARS[int(o) & 15]}" for o in ops)


def align(ref: bytes, alt: bytes, params: Tuple[int, int, int, int], strategy: int) -> Tuple[str, int]:
    """Oracle alignment -> (cigar text, offset)."""
    ops, res = align_ops(ref, alt, params, strategy)
    return ops_to_text(ops), res.offset


def align_ops(ref: bytes, alt: bytes, params, strategy) -> Tuple[List[int], OracleResult]:
    lib = _load_oracle()
    cap = len(ref) + len(alt) + 8
    ops = (C.c_uint32 * cap)()
    res = OracleResult()
    rc = lib.oracle_sw_align(ref, len(ref), alt, len(alt), params[0], params[1], params[2], params[3], strategy,
                             ops, cap, C.byref(res))
    if rc != 0:
        raise RuntimeError(f"oracle_sw_align rc={rc}")
    return list(ops[: res.n_ops]), res


def score3(ref: bytes, alt: bytes, params, strategy) -> Tuple[int, int, int]:
    """Oracle (score, row, column) of the cell the traceback starts from (sw.cpp:155-170)."""
    _, res = align_ops(ref, alt, params, strategy)
    return res.score, res.start_i, res.start_j


def path_score(cigar: str, offset: int, ref: bytes, alt: bytes, params) -> int:
    """Score of the alignment a CIGAR describes, under the recurrence of sw.cpp:55-93: per-base match / mismatch for M,
    -open - (len - 1) * extend per I / D run; S clips cost nothing.  For the INDEL strategies this is H[tl][ql] resp.
    H[mqe_t][ql], which ties the score-only result to the reference's OUTPUT where its ScoreMax has no field."""
    import re
    m, x, go, ge = abs(params[0]), -abs(params[1]), abs(params[2]), abs(params[3])
    i, j, s = offset, 0, 0
    for ln, op in re.findall(r"(\d+)([MIDS])", cigar):
        ln = int(ln)
        if op == "M":
            s += sum(m if ref[i + k] == alt[j + k] else x for k in range(ln))
            i += ln
            j += ln
        elif op == "I":
            s -= go + (ln - 1) * ge
            j += ln
        elif op == "D":
            s -= go + (ln - 1) * ge
            i += ln
        else:
            j += ln
    return s


def ref_scoremax(ref: bytes, alt: bytes, params, strategy) -> dict:
    """The reference's own ScoreMax (its calculateMatrix, oracle/_ref)."""
    lib = _load_ref()
    out = (C.c_int32 * 6)()
    if lib.mgl_ref_scoremax(ref, len(ref), alt, len(alt), params[0], params[1], params[2], params[3], strategy, out) != 0:
        raise MemoryError("mgl_ref_scoremax")
    return dict(zip(("mqe", "mqe_t", "max", "max_t", "max_q", "seg_length"), list(out)))


REF_SCALAR, REF_AVX, REF_DISPATCH = 0, 1, 2


def ref_align(ref: bytes, alt: bytes, params, strategy, kind: int = REF_DISPATCH) -> Tuple[str, int]:
    """The reference's own C++ (oracle/_ref) -> (cigar text, offset)."""
    lib = _load_ref()
    cap = 12 * (len(ref) + len(alt)) + 64
    buf = C.create_string_buffer(cap)
    off = C.c_int32()
    n = lib.mgl_ref_align(kind, ref, len(ref), alt, len(alt), params[0], params[1], params[2], params[3], strategy,
                          buf, cap, C.byref(off))
    if n < 0:
        raise RuntimeError("cigar buffer too small")
    return buf.value.decode(), off.value


def ref_align_jni(ref: bytes, alt: bytes, params, strategy) -> Tuple[str, int]:
    """Through the reference's exported JNI symbol (readsBuffer = ref||alt, zeroed cigar buffer, trimmed)."""
    lib = _load_ref()
    reads = ref + alt
    cap = 12 * (len(ref) + len(alt)) + 64
    buf = C.create_string_buffer(cap)
    off = lib.mgl_ref_align_jni(reads, len(ref), len(alt), params[0], params[1], params[2], params[3], strategy, buf)
    return buf.value.decode().strip(), off


def ref_max_threads() -> int:
    """Host threads the reference runs on: the CPUs this process may use.  NOT omp_get_max_threads(): launchers
    such as torchrun export OMP_NUM_THREADS=1, which would silently time the reference on one core."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def ref_align_batch(arena: np.ndarray, ref_off, alt_off, ref_len, alt_len, params, strategy: int = SOFTCLIP,
                    strategies: Optional[np.ndarray] = None, threads: int = 1, kind: int = REF_DISPATCH,
                    want_output: bool = True, cigar_stride: int = 0):
    """Run the reference over a batch; returns (seconds, offsets, [cigar strings])."""
    lib = _load_ref()
    n = len(ref_len)
    arena = np.ascontiguousarray(arena, dtype=np.uint8)
    ref_off = np.ascontiguousarray(ref_off, dtype=np.uint64)
    alt_off = np.ascontiguousarray(alt_off, dtype=np.uint64)
    ref_len = np.ascontiguousarray(ref_len, dtype=np.uint32)
    alt_len = np.ascontiguousarray(alt_len, dtype=np.uint32)
    st_ptr = None
    if strategies is not None:
        strategies = np.ascontiguousarray(strategies, dtype=np.uint8)
        st_ptr = strategies.ctypes.data
    offsets = cig = None
    if want_output:
        if cigar_stride <= 0:
            cigar_stride = int(4 * (int(ref_len.max()) + int(alt_len.max())) + 16) if n else 16
        offsets = np.zeros(n, dtype=np.int32)
        cig = np.zeros(n * cigar_stride, dtype=np.uint8)
    secs = lib.mgl_ref_align_batch(kind, n, arena.ctypes.data, ref_off.ctypes.data, alt_off.ctypes.data,
                                   ref_len.ctypes.data, alt_len.ctypes.data, params[0], params[1], params[2],
                                   params[3], strategy, st_ptr, threads,
                                   offsets.ctypes.data if want_output else None,
                                   cig.ctypes.data if want_output else None, cigar_stride if want_output else 0)
    if not want_output:
        return secs, None, None
    raw = cig.reshape(n, cigar_stride)
    texts = [bytes(row).split(b"\0", 1)[0].decode() for row in raw]
    return secs, offsets, texts


def ref_check_batch(arena: np.ndarray, ref_off, alt_off, ref_len, alt_len, params, strategies, got_offsets, cigar_start,
                    ops, threads: Optional[int] = None, kind: int = REF_DISPATCH, max_bad: int = 8):
    """Every pair through the reference (one pair per OpenMP task) and compared in place with the caller's results:
    offsets[k] and the packed elements ops[cigar_start[k]:cigar_start[k+1]] printed as sw.cpp:251-252 prints them.
    Returns (differing pairs, their first indices, seconds).  Full-size parity without n x stride bytes of text."""
    lib = _load_ref()
    n = len(ref_len)
    arena = np.ascontiguousarray(arena, dtype=np.uint8)
    ref_off = np.ascontiguousarray(ref_off, dtype=np.uint64)
    alt_off = np.ascontiguousarray(alt_off, dtype=np.uint64)
    ref_len = np.ascontiguousarray(ref_len, dtype=np.uint32)
    alt_len = np.ascontiguousarray(alt_len, dtype=np.uint32)
    strategies = np.ascontiguousarray(strategies, dtype=np.int32)
    got_offsets = np.ascontiguousarray(got_offsets, dtype=np.int32)
    cigar_start = np.ascontiguousarray(cigar_start, dtype=np.uint64)
    ops = np.ascontiguousarray(ops, dtype=np.uint32)
    assert len(strategies) == n and len(got_offsets) == n and len(cigar_start) == n + 1 and int(cigar_start[-1]) <= len(ops)
    bad = np.full(max_bad, -1, dtype=np.int64)
    secs = C.c_double()
    nbad = lib.mgl_ref_check_batch(kind, n, arena.ctypes.data, ref_off.ctypes.data, alt_off.ctypes.data, ref_len.ctypes.data,
                                   alt_len.ctypes.data, params[0], params[1], params[2], params[3], SOFTCLIP,
                                   strategies.ctypes.data, threads or ref_max_threads(), got_offsets.ctypes.data,
                                   cigar_start.ctypes.data, ops.ctypes.data, bad.ctypes.data, max_bad, C.byref(secs))
    return int(nbad), [int(x) for x in bad if x >= 0], secs.value

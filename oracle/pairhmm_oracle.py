"""TEST INFRASTRUCTURE: Python access to the PairHMM oracle (oracle/pairhmm_oracle.c, a plain-C restatement of
the reference's forward algorithm, compute_prob_scalar.cc + ...MicrosoftPairHmm.cc:77-222) and, when it was built
here, to the reference's own PairHMM (oracle/_ref/libmgl_pairhmm_ref.so).  Only tests/, __graft_entry__.smoke()
and the cpu_baseline / --impl reference legs of the benches may import this."""
import ctypes as C
import os
import subprocess
from typing import List, Sequence, Tuple

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "libpairhmm_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libmgl_pairhmm_ref.so")
_oracle = None
_ref = None

Read = Tuple[bytes, bytes, bytes, bytes, bytes]      # bases, base quals, insertion GOP, deletion GOP, GCP (phred, not +33)


def build(force: bool = False) -> None:
    if force or not os.path.exists(ORACLE_SO):
        subprocess.check_call(["make", "-C", HERE, ORACLE_SO])


def pack(reads: Sequence[Read], haps: Sequence[bytes]):
    """-> (lengths int32, reads uint8, haps uint8): the three direct buffers of computeLikelihoodsNative (.cc:84-122)"""
    lengths = np.array([len(reads)] + [len(r[0]) for r in reads] + [len(haps)] + [len(h) for h in haps], dtype=np.int32)
    for r in reads:
        assert all(len(x) == len(r[0]) for x in r)
    rb = np.frombuffer(b"".join(b"".join(r) for r in reads), dtype=np.uint8).copy() if reads else np.zeros(0, np.uint8)
    hb = np.frombuffer(b"".join(haps), dtype=np.uint8).copy() if haps else np.zeros(0, np.uint8)
    return lengths, rb, hb


def _load_oracle():
    global _oracle
    if _oracle is None:
        build()
        lib = C.CDLL(ORACLE_SO)
        lib.oracle_pairhmm.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        lib.oracle_pairhmm.restype = C.c_int
        _oracle = lib
    return _oracle


def likelihoods(reads: Sequence[Read], haps: Sequence[bytes], use_double: bool = False):
    """Oracle: (log10 likelihoods [n_reads, n_haps], which pairs were recomputed in double)."""
    lengths, rb, hb = pack(reads, haps)
    out = np.zeros((len(reads), len(haps)), dtype=np.float64)
    esc = np.zeros((len(reads), len(haps)), dtype=np.uint8)
    rc = _load_oracle().oracle_pairhmm(lengths.ctypes.data, rb.ctypes.data, hb.ctypes.data, int(use_double), out.ctypes.data, esc.ctypes.data)
    if rc != 0:
        raise RuntimeError(f"oracle_pairhmm rc={rc}")
    return out, esc


def have_ref() -> bool:
    return os.path.exists(REF_SO)


def _load_ref():
    global _ref
    if _ref is None:
        lib = C.CDLL(REF_SO)
        lib.mgl_ref_pairhmm.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        lib.mgl_ref_pairhmm.restype = None
        lib.mgl_ref_pairhmm_threads.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int]
        lib.mgl_ref_pairhmm_threads.restype = C.c_double
        lib.mgl_ref_pairhmm_max_threads.restype = C.c_int
        _ref = lib
    return _ref


def ref_likelihoods(reads: Sequence[Read], haps: Sequence[bytes], use_double: bool = False) -> np.ndarray:
    """The reference's own JNI entry (AVX2 float with escalation to double, or double throughout)."""
    lengths, rb, hb = pack(reads, haps)
    out = np.zeros((len(reads), len(haps)), dtype=np.float64)
    _load_ref().mgl_ref_pairhmm(lengths.ctypes.data, rb.ctypes.data, hb.ctypes.data, int(use_double), out.ctypes.data)
    return out


def ref_max_threads() -> int:
    return int(_load_ref().mgl_ref_pairhmm_max_threads())


def ref_likelihoods_timed(lengths, rb, hb, use_double=False, threads=0, reads_per_call=16):
    """-> (seconds, out): packed buffers in, the reads spread over OpenMP threads (one JNI call per slice of reads)."""
    R = int(lengths[0]); H = int(lengths[1 + R])
    out = np.zeros((R, H), dtype=np.float64)
    secs = _load_ref().mgl_ref_pairhmm_threads(lengths.ctypes.data, rb.ctypes.data, hb.ctypes.data, int(use_double), out.ctypes.data,
                                               threads or ref_max_threads(), reads_per_call)
    return float(secs), out


def normalize(quals: bytes, floor: int = 0) -> bytes:
    """MicrosoftPairHmmUnitTest.java:131-137: phred+33 text -> phred, clamped below"""
    return bytes(max(q - 33, floor) for q in quals)


def parse_testdata(text: str) -> List[Tuple[bytes, Read, float]]:
    """lines of pairhmm-testdata.txt: hap read quals insGOP delGOP GCP expected (MicrosoftPairHmmUnitTest.java:84-101)"""
    cases = []
    for line in text.splitlines():
        if not line.strip() or line.startswith("#"):
            continue
        hap, rs, q, i, d, c, exp = line.split()
        read = (rs.encode(), normalize(q.encode(), 6), normalize(i.encode()), normalize(d.encode()), normalize(c.encode()))
        cases.append((hap.encode(), read, float(exp)))
    return cases

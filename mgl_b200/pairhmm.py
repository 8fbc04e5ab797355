"""Host-side mirror of the reference's Java class MicrosoftPairHmm (src/main/java/com/microsoft/mgl/pairhmm/
MicrosoftPairHmm.java:20-125) over the C-ABI of libmgl_pairhmm.so (include/mgl_pairhmm.h): load(), initialize(args),
computeLikelihoods(readDataArray, haplotypeDataArray, likelihoodArray), done() with the reference's argument meaning
-- every read against every haplotype, log10 likelihoods read-major.  There is no CPU fallback: without the built
library or a B200 every call raises."""
import ctypes as C
import os
from dataclasses import dataclass
from typing import Optional, Sequence

import numpy as np

LIB_PATH = os.environ.get("MGL_PAIRHMM_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "libmgl_pairhmm.so")
OK, E_INVALID, E_NO_DEVICE, E_CUDA, E_TOO_LONG, E_TOO_LARGE = 0, -1, -2, -3, -4, -5
NOALLOC_MAX_READS, NOALLOC_MAX_HAPS, NOALLOC_MAX_PAIRS = 4096, 1024, 65536       # include/mgl_pairhmm.h
NOALLOC_MAX_READ_BASES = NOALLOC_MAX_HAP_BASES = 262144
NOALLOC_NO_SYNC = 1
_lib = None


class MglPairHmmError(RuntimeError):
    def __init__(self, code, msg=""):
        super().__init__(f"mgl_pairhmm error {code}: {msg}")
        self.code = code


class Stats(C.Structure):
    _fields_ = [(k, C.c_uint64) for k in ("pairs", "cells", "pairs_double", "launches", "h2d_bytes", "d2h_bytes")] + [("device_ms", C.c_double)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MglPairHmmError(E_NO_DEVICE, f"{LIB_PATH} not built (python -c 'import __graft_entry__ as g; g.build()')")
        L = C.CDLL(LIB_PATH)
        L.mgl_pairhmm_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
        L.mgl_pairhmm_destroy.argtypes = [C.c_void_p]
        L.mgl_pairhmm_destroy.restype = None
        L.mgl_pairhmm_last_error.argtypes = [C.c_void_p]
        L.mgl_pairhmm_last_error.restype = C.c_char_p
        L.mgl_pairhmm_compute.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.POINTER(Stats)]
        L.mgl_pairhmm_resident_prepare.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.mgl_pairhmm_resident_run.argtypes = [C.c_void_p, C.c_int, C.POINTER(Stats)]
        L.mgl_pairhmm_resident_fetch.argtypes = [C.c_void_p, C.c_void_p]
        L.mgl_pairhmm_compute_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        L.mgl_pairhmm_noalloc_scratch_bytes.restype = C.c_uint64
        L.mgl_pairhmm_noalloc_last_error.restype = C.c_char_p
        _lib = L
    return _lib


@dataclass
class ReadDataHolder:                 # org.broadinstitute.gatk.nativebindings.pairhmm.ReadDataHolder
    readBases: bytes
    readQuals: bytes
    insertionGOP: bytes
    deletionGOP: bytes
    overallGCP: bytes


@dataclass
class HaplotypeDataHolder:            # ...pairhmm.HaplotypeDataHolder
    haplotypeBases: bytes


@dataclass
class PairHMMNativeArguments:         # ...pairhmm.PairHMMNativeArguments
    useDoublePrecision: bool = False
    maxNumberOfThreads: int = 1


def pack(reads: Sequence[ReadDataHolder], haps: Sequence[HaplotypeDataHolder]):
    """the three input buffers exactly as MicrosoftPairHmm.java:60-98 fills them"""
    lengths = np.array([len(reads)] + [len(r.readBases) for r in reads] + [len(haps)] + [len(h.haplotypeBases) for h in haps], dtype=np.int32)
    rb = np.frombuffer(b"".join(r.readBases + r.readQuals + r.insertionGOP + r.deletionGOP + r.overallGCP for r in reads), dtype=np.uint8).copy() \
        if reads else np.zeros(0, np.uint8)
    hb = np.frombuffer(b"".join(h.haplotypeBases for h in haps), dtype=np.uint8).copy() if haps else np.zeros(0, np.uint8)
    return lengths, rb, hb


def compute_noalloc(lengths: np.ndarray, reads, haps, out, use_double: bool = False, device: int = 0, stream: int = 0, flags: int = 0):
    """mgl_pairhmm_compute_noalloc (no counterpart in the reference): the job out of the library's static device
    array, no handle and no allocation.  `lengths` is a host int32 array; `reads`, `haps`, `out` are numpy arrays or
    integer addresses (device or host pointers); `stream` is a cudaStream_t as an integer, 0 = the default stream."""
    def addr(x):
        return x.ctypes.data if isinstance(x, np.ndarray) else int(x)
    rc = lib().mgl_pairhmm_compute_noalloc(device, lengths.ctypes.data, addr(reads), addr(haps), int(bool(use_double)), addr(out), stream or None, flags)
    if rc != OK:
        raise MglPairHmmError(rc, lib().mgl_pairhmm_noalloc_last_error().decode(errors="replace"))
    return out


class MicrosoftPairHmm:
    def __init__(self, device: int = 0):
        self._device = device
        self._h = None
        self._use_double = False
        self.last_stats: dict = {}

    def load(self, temp_dir=None) -> bool:
        """True when the native library is present and a B200 can be opened (.java:34-36)"""
        try:
            self._ensure()
            return True
        except MglPairHmmError:
            return False

    def _ensure(self):
        if self._h is None:
            h = C.c_void_p()
            rc = lib().mgl_pairhmm_create(self._device, C.byref(h))
            if rc != OK:
                raise MglPairHmmError(rc, "cannot open the GPU")
            self._h = h
        return self._h

    def initialize(self, args: Optional[PairHMMNativeArguments] = None):
        args = args or PairHMMNativeArguments()
        self._use_double = bool(args.useDoublePrecision)

    def _err(self, rc):
        return MglPairHmmError(rc, lib().mgl_pairhmm_last_error(self._h).decode(errors="replace"))

    def computeLikelihoods(self, readDataArray, haplotypeDataArray, likelihoodArray: Optional[np.ndarray] = None) -> np.ndarray:
        lengths, rb, hb = pack(readDataArray, haplotypeDataArray)
        return self.compute_buffers(lengths, rb, hb, likelihoodArray).reshape(len(readDataArray), len(haplotypeDataArray))

    def compute_buffers(self, lengths, reads, haps, out: Optional[np.ndarray] = None, use_double: Optional[bool] = None) -> np.ndarray:
        """what a JNI / FFI caller passes: the packed buffers of computeLikelihoodsNative"""
        R = int(lengths[0]); H = int(lengths[1 + R])
        if out is None:
            out = np.zeros(R * H, dtype=np.float64)
        st = Stats()
        ud = self._use_double if use_double is None else bool(use_double)
        rc = lib().mgl_pairhmm_compute(self._ensure(), lengths.ctypes.data, reads.ctypes.data, haps.ctypes.data, int(ud), out.ctypes.data, C.byref(st))
        if rc != OK:
            raise self._err(rc)
        self.last_stats = st.as_dict()
        return out

    # device-resident timing path used by tools/bench_pairhmm.py
    def resident_prepare(self, lengths, reads, haps):
        rc = lib().mgl_pairhmm_resident_prepare(self._ensure(), lengths.ctypes.data, reads.ctypes.data, haps.ctypes.data)
        if rc != OK:
            raise self._err(rc)
        self._resident = (int(lengths[0]), int(lengths[1 + int(lengths[0])]))

    def resident_run(self, use_double=False) -> dict:
        st = Stats()
        rc = lib().mgl_pairhmm_resident_run(self._ensure(), int(use_double), C.byref(st))
        if rc != OK:
            raise self._err(rc)
        return st.as_dict()

    def resident_fetch(self) -> np.ndarray:
        R, H = self._resident
        out = np.zeros(R * H, dtype=np.float64)
        rc = lib().mgl_pairhmm_resident_fetch(self._ensure(), out.ctypes.data)
        if rc != OK:
            raise self._err(rc)
        return out.reshape(R, H)

    def done(self):
        if self._h is not None:
            lib().mgl_pairhmm_destroy(self._h)
            self._h = None

    close = done

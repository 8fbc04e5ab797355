"""GPU (B200): the CUDA PairHMM, called through the C-ABI (libmgl_pairhmm.so), against the reference's own fixture
values (pairhmm-testdata.txt), the reference-generated batches and the oracle on fresh seeded inputs.  Tolerance
1e-5 on the log10 likelihood -- the bar of the reference's own test (MicrosoftPairHmmUnitTest.java:101)."""
import ctypes as C
import gzip
import json
import os
import random

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOL = 1e-5


@pytest.fixture(scope="module")
def ph_golden():
    with gzip.open(os.path.join(ROOT, "tests", "golden", "pairhmm_golden.json.gz"), "rb") as f:
        return json.loads(f.read().decode())


@pytest.fixture(scope="module")
def hmm():
    from mgl_b200 import pairhmm
    h = pairhmm.MicrosoftPairHmm(0)
    h._ensure()                      # raises when libmgl_pairhmm.so or the B200 is missing: nothing to fall back to
    yield h
    h.done()


@pytest.fixture(scope="module")
def oracle_ph():
    from oracle import pairhmm_oracle
    pairhmm_oracle.build()
    return pairhmm_oracle


def _holders(reads, haps):
    from mgl_b200 import pairhmm
    return ([pairhmm.ReadDataHolder(*r) for r in reads], [pairhmm.HaplotypeDataHolder(h) for h in haps])


def _read(r):
    return tuple(x.encode("latin1") for x in r)


def test_reference_fixture_values(hmm, ph_golden):
    from mgl_b200 import pairhmm
    for use_double in (False, True):
        hmm.initialize(pairhmm.PairHMMNativeArguments(useDoublePrecision=use_double))
        for c in ph_golden["file"]:
            r, h = _holders([_read(c["read"])], [c["hap"].encode("latin1")])
            got = hmm.computeLikelihoods(r, h)[0, 0]
            assert abs(got - c["expected_file"]) <= TOL, (use_double, got, c["expected_file"])
    hmm.initialize(None)
    s = ph_golden["simple"]
    r, h = _holders([_read(s["read"])], [s["hap"].encode()])
    assert abs(hmm.computeLikelihoods(r, h)[0, 0] - s["expected"]) <= TOL


def test_reference_generated_batches(hmm, ph_golden):
    from mgl_b200 import pairhmm
    n = 0
    redone = 0
    for b in ph_golden["batch"]:
        r, h = _holders([_read(x) for x in b["reads"]], [x.encode("latin1") for x in b["haps"]])
        keep = np.array(b["finite"], dtype=bool)
        for use_double, key in ((False, "ref_float"), (True, "ref_double")):
            hmm.initialize(pairhmm.PairHMMNativeArguments(useDoublePrecision=use_double))
            got = hmm.computeLikelihoods(r, h)
            exp = np.array([[v if v is not None else np.nan for v in row] for row in b[key]])
            assert np.all(np.abs(got - exp)[keep] <= TOL), (key, float(np.nanmax(np.abs(got - exp)[keep])))
            if not use_double:
                redone += hmm.last_stats["pairs_double"]
        n += int(keep.sum())
    hmm.initialize(None)
    assert n > 1000 and redone > 100        # the float -> double escalation ran on the device


@pytest.mark.parametrize("max_read", [30, 60, 90, 120, 150, 190, 250, 700])
def test_against_oracle_every_band_shape(hmm, oracle_ph, max_read):
    """rows per lane 1..6 or 8 are picked from the longest read; 700-base reads take three bands through the
    shared-memory carry row; reads of one base, haplotypes of one base, 'N' on both sides"""
    rng = random.Random(max_read)
    haps = [bytes(rng.choice(b"ACGTN") for _ in range(n)) for n in (1, 2, 9, 33, 150, 320, 511)]
    reads = []
    for k in range(40):
        n = max_read if k == 0 else 1 if k == 1 else rng.randint(1, max_read)
        if k % 3 == 0:
            src = haps[-1] * 3
            s = rng.randrange(len(src) - n)
            b = bytearray(src[s:s + n])
            for _ in range(rng.randint(0, 3)):
                b[rng.randrange(n)] = rng.choice(b"ACGTN")
            b = bytes(b)
        else:
            b = bytes(rng.choice(b"ACGTN") for _ in range(n))
        reads.append((b, bytes(rng.randint(6, 45) for _ in range(n)), bytes(rng.randint(10, 50) for _ in range(n)),
                      bytes(rng.randint(10, 50) for _ in range(n)), bytes(rng.randint(5, 20) for _ in range(n))))
    from mgl_b200 import pairhmm
    r, h = _holders(reads, haps)
    for use_double in (False, True):
        hmm.initialize(pairhmm.PairHMMNativeArguments(useDoublePrecision=use_double))
        got = hmm.computeLikelihoods(r, h)
        exp, esc = oracle_ph.likelihoods(reads, haps, use_double)
        ok = np.isfinite(exp)                        # a 700-base unrelated read underflows even double: -inf on both sides
        assert ok.mean() > 0.6 and np.all(np.abs(got[ok] - exp[ok]) <= TOL), float(np.abs(got[ok] - exp[ok]).max())
        assert np.all(~np.isfinite(got[~ok]))
        if not use_double:
            # which pairs escalate is decided by a float value next to the 1e-28 threshold: allow the few borderline ones
            assert abs(int(hmm.last_stats["pairs_double"]) - int(esc.sum())) <= 2
    hmm.initialize(None)


def test_jni_symbols(hmm, ph_golden):
    """Java_..._MicrosoftPairHmm_initNative / computeLikelihoodsNative through a table-backed JNIEnv
    (GetDirectBufferAddress = identity): the buffers MicrosoftPairHmm.java:60-103 builds"""
    from mgl_b200 import pairhmm
    L = pairhmm.lib()
    GETADDR = C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_void_p)
    ident = GETADDR(lambda env, buf: buf)
    table = (C.c_void_p * 235)()
    table[230] = C.cast(ident, C.c_void_p)
    env = C.pointer(C.c_void_p(C.addressof(table)))
    init = L.Java_com_microsoft_mgl_pairhmm_MicrosoftPairHmm_initNative
    init.argtypes = [C.c_void_p, C.c_void_p, C.c_uint8, C.c_int]
    init.restype = None
    fn = L.Java_com_microsoft_mgl_pairhmm_MicrosoftPairHmm_computeLikelihoodsNative
    fn.argtypes = [C.c_void_p] * 6
    fn.restype = None
    b = ph_golden["batch"][0]
    r, h = _holders([_read(x) for x in b["reads"]], [x.encode("latin1") for x in b["haps"]])
    lengths, rb, hb = pairhmm.pack(r, h)
    keep = np.array(b["finite"], dtype=bool)
    for use_double, key in ((0, "ref_float"), (1, "ref_double")):
        out = np.zeros(len(r) * len(h), dtype=np.float64)
        init(env, None, use_double, 1)
        fn(env, None, lengths.ctypes.data, rb.ctypes.data, hb.ctypes.data, out.ctypes.data)
        exp = np.array([[v if v is not None else np.nan for v in row] for row in b[key]])
        assert np.all(np.abs(out.reshape(exp.shape) - exp)[keep] <= TOL)
    init(env, None, 0, 1)


def test_bad_arguments(hmm):
    from mgl_b200 import pairhmm
    r, h = _holders([(b"", b"", b"", b"", b"")], [b"ACGT"])
    with pytest.raises(pairhmm.MglPairHmmError):
        hmm.computeLikelihoods(r, h)
    r, h = _holders([(b"A", b"!", b"!", b"!", b"!")], [b""])
    with pytest.raises(pairhmm.MglPairHmmError):
        hmm.computeLikelihoods(r, h)
    assert hmm.computeLikelihoods([], []).size == 0

"""GPU (B200): mgl_pairhmm_compute_noalloc -- the PairHMM job out of the library's static __device__ array, for callers
that cannot call cudaMalloc (no counterpart in the reference; include/mgl_pairhmm.h) -- called through the C-ABI and
checked against the existing entry point mgl_pairhmm_compute at the tolerance of the other PairHMM tests (1e-5 on the
log10 likelihood), on calls in a row that share the array.  Collected last on purpose (pytest -x)."""
import ctypes as C
import random

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-5


def _job(seed, n_reads, hap_lens, max_read):
    """reads cut from the last haplotype with a few edits (likely), and unrelated reads at modest qualities whose float
    likelihood falls below 1e-28 -- the pairs the float pass hands to the double pass through the redo list"""
    from mgl_b200 import pairhmm
    rng = random.Random(seed)
    haps = [bytes(rng.choice(b"ACGTN") for _ in range(n)) for n in hap_lens]
    reads = []
    for k in range(n_reads):
        n = max_read if k == 0 else 1 if k == 1 else rng.randint(max(1, max_read // 3), max_read)
        if k % 2 == 0:
            src = haps[-1] * 3
            s = rng.randrange(len(src) - n)
            b = bytearray(src[s:s + n])
            for _ in range(rng.randint(0, 3)):
                b[rng.randrange(n)] = rng.choice(b"ACGT")
            b = bytes(b)
        else:
            b = bytes(rng.choice(b"ACGT") for _ in range(n))
        reads.append(pairhmm.ReadDataHolder(b, bytes(rng.randint(15, 30) for _ in range(n)), bytes(rng.randint(10, 50) for _ in range(n)),
                                            bytes(rng.randint(10, 50) for _ in range(n)), bytes(rng.randint(5, 20) for _ in range(n))))
    return pairhmm.pack(reads, [pairhmm.HaplotypeDataHolder(h) for h in haps])


def _same(got, exp):
    fin = np.isfinite(exp)
    assert np.array_equal(fin, np.isfinite(got))
    assert np.array_equal(got[~fin], exp[~fin])                  # -inf on both sides where even double underflows
    assert fin.mean() > 0.5 and np.all(np.abs(got[fin] - exp[fin]) <= TOL), float(np.abs(got[fin] - exp[fin]).max())


@pytest.fixture(scope="module")
def hmm():
    from mgl_b200 import pairhmm
    h = pairhmm.MicrosoftPairHmm(0)
    h._ensure()                      # raises when libmgl_pairhmm.so or the B200 is missing: nothing to fall back to
    yield h
    h.done()


def test_noalloc_matches_compute_on_calls_in_a_row(hmm):
    from mgl_b200 import pairhmm
    assert 0 < pairhmm.lib().mgl_pairhmm_noalloc_scratch_bytes() <= 4 << 20
    # a large job, then a smaller one (whatever the first left in the shared array must not leak into the second),
    # then the first again; 150-base reads run two pairs per warp, 250 one pair per warp, 600 three bands
    jobs = [_job(1, 96, (1, 9, 150, 320, 511), 150), _job(2, 24, (33, 280), 250), _job(3, 12, (40, 300, 450), 600), _job(1, 96, (1, 9, 150, 320, 511), 150)]
    escalated = 0
    for lengths, rb, hb in jobs:
        R = int(lengths[0]); H = int(lengths[1 + R])
        for use_double in (False, True):
            exp = hmm.compute_buffers(lengths, rb, hb, use_double=use_double)
            if not use_double:
                escalated += hmm.last_stats["pairs_double"]
            got = np.full(R * H, 12345.0)
            pairhmm.compute_noalloc(lengths, rb, hb, got, use_double=use_double)
            _same(got, exp)
            again = np.full(R * H, -1.0)                          # twice in a row on the same inputs
            pairhmm.compute_noalloc(lengths, rb, hb, again, use_double=use_double)
            assert np.array_equal(got, again)
    assert escalated > 50            # the redo list (the scratch this variant keeps in the static array) was really used


def test_noalloc_device_buffers_and_stream(hmm):
    """the caller's buffers on the device, a caller's stream, no synchronisation inside the call"""
    import torch
    from mgl_b200 import pairhmm
    lengths, rb, hb = _job(7, 40, (120, 410), 150)
    R = int(lengths[0]); H = int(lengths[1 + R])
    exp = hmm.compute_buffers(lengths, rb, hb, use_double=False)
    d_reads = torch.from_numpy(rb).cuda()
    d_haps = torch.from_numpy(hb).cuda()
    d_out = torch.zeros(R * H, dtype=torch.float64, device="cuda")
    st = torch.cuda.Stream()
    torch.cuda.synchronize()
    for _ in range(2):
        d_out.fill_(7.0)
        torch.cuda.synchronize()
        pairhmm.compute_noalloc(lengths, d_reads.data_ptr(), d_haps.data_ptr(), d_out.data_ptr(), stream=st.cuda_stream, flags=pairhmm.NOALLOC_NO_SYNC)
        st.synchronize()
        _same(d_out.cpu().numpy(), exp)


def test_noalloc_rejects_what_does_not_fit(hmm):
    from mgl_b200 import pairhmm
    L = pairhmm.lib()
    out = np.zeros(8)
    one = np.frombuffer(b"A!!!!", dtype=np.uint8).copy()
    # more pairs than the redo list holds: 300 reads x 300 haplotypes
    R = H = 300
    lengths = np.array([R] + [1] * R + [H] + [1] * H, dtype=np.int32)
    rb = np.tile(one, R)
    hb = np.frombuffer(b"A" * H, dtype=np.uint8).copy()
    rc = L.mgl_pairhmm_compute_noalloc(0, lengths.ctypes.data, rb.ctypes.data, hb.ctypes.data, 0, out.ctypes.data, None, 0)
    assert rc == pairhmm.E_TOO_LARGE and b"MGL_PH_NOALLOC_MAX_PAIRS" in L.mgl_pairhmm_noalloc_last_error()
    # more reads / more haplotype bases than the limits
    lengths = np.array([pairhmm.NOALLOC_MAX_READS + 1] + [1] * (pairhmm.NOALLOC_MAX_READS + 1) + [1, 1], dtype=np.int32)
    rc = L.mgl_pairhmm_compute_noalloc(0, lengths.ctypes.data, rb.ctypes.data, hb.ctypes.data, 0, out.ctypes.data, None, 0)
    assert rc == pairhmm.E_TOO_LARGE
    lengths = np.array([1, 1, 2, pairhmm.NOALLOC_MAX_HAP_BASES, 1], dtype=np.int32)
    rc = L.mgl_pairhmm_compute_noalloc(0, lengths.ctypes.data, rb.ctypes.data, hb.ctypes.data, 0, out.ctypes.data, None, 0)
    assert rc == pairhmm.E_TOO_LARGE and b"MGL_PH_NOALLOC_MAX_HAP_BASES" in L.mgl_pairhmm_noalloc_last_error()
    with pytest.raises(pairhmm.MglPairHmmError):
        pairhmm.compute_noalloc(np.array([1, 0, 1, 4], dtype=np.int32), one, hb, out)      # an empty read
    # and a job exactly at the pair limit still runs
    R, H = 256, 256
    lengths = np.array([R] + [1] * R + [H] + [1] * H, dtype=np.int32)
    got = np.zeros(R * H)
    pairhmm.compute_noalloc(lengths, np.tile(one, R), np.frombuffer(b"A" * H, dtype=np.uint8).copy(), got)
    exp = hmm.compute_buffers(lengths, np.tile(one, R), np.frombuffer(b"A" * H, dtype=np.uint8).copy(), use_double=False)
    _same(got, exp)

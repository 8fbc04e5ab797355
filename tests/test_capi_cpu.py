"""CPU: the C-ABI library builds, loads, exports every symbol include/mgl_sw.h
declares, plans batches, and refuses to compute without a GPU (no fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as g
    from mgl_b200 import native
    if not os.path.exists(native.LIB_PATH):
        g.build()
    return native.lib()


def test_exports_every_declared_symbol(lib):
    hdr = open(os.path.join(ROOT, "include", "mgl_sw.h")).read()
    names = re.findall(r"MGL_SW_API\s+[\w\s\*]+?\b(mgl_sw_\w+)\s*\(", hdr)
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), n
    for n in ("initNative", "alignNative", "doneNative"):      # ...MicrosoftSmithWaterman.h:15,23,31
        assert hasattr(lib, "Java_com_microsoft_mgl_smithwaterman_MicrosoftSmithWaterman_" + n)


def test_synth_library_exports_its_header():
    from mgl_b200 import synth
    hdr = open(os.path.join(ROOT, "include", "mgl_synth.h")).read()
    names = re.findall(r"MGL_SYNTH_API\s+[\w\s\*]+?\b(mgl_sw_synth_\w+)\s*\(", hdr)
    assert sorted(names) == ["mgl_sw_synth_fill", "mgl_sw_synth_sizes"]
    for n in names:
        assert hasattr(synth.lib(), n), n


def test_no_gpu_no_handle(lib):
    """no device in this container: every way of getting a handle refuses (there is no CPU path to fall back to)"""
    from mgl_b200 import native
    if lib.mgl_sw_device_count() > 0:
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    assert lib.mgl_sw_create(0, C.byref(h)) == native.E_NO_DEVICE and not h.value
    devs = (C.c_int * 2)(0, 1)
    assert lib.mgl_sw_create_multi(devs, 2, C.byref(h)) == native.E_NO_DEVICE and not h.value
    assert lib.mgl_sw_create_multi(None, 0, C.byref(h)) == native.E_INVALID
    assert lib.mgl_sw_default_handle(C.byref(h)) == native.E_NO_DEVICE and not h.value
    lib.mgl_sw_shutdown()                                   # nothing to release: returns at once
    buf, ln, off = C.create_string_buffer(64), C.c_int32(), C.c_int32()
    assert lib.mgl_sw_align_one(b"ACGT", 4, b"ACGT", 4, 10, -15, -30, -5, 1, buf, 64, C.byref(ln), C.byref(off)) == native.E_NO_DEVICE


def test_format_cigar(lib):
    ops = np.array([(3 << 4) | 0, (2 << 4) | 2, (15 << 4) | 0, (4 << 4) | 4, (1 << 4) | 1], dtype=np.uint32)
    buf = C.create_string_buffer(64)
    n = lib.mgl_sw_format_cigar(ops.ctypes.data, len(ops), buf, 64)
    assert buf.value == b"3M2D15M4S1I" and n == 11
    assert lib.mgl_sw_format_cigar(ops.ctypes.data, len(ops), buf, 4) == -1


def _pack2_numpy(a):
    """sw_pack.h restated: code = (byte >> 1) & 3, base k in bits 2 (k % 4) of byte k / 4, whole u32 words."""
    out = np.zeros(((len(a) + 15) // 16) * 4, dtype=np.uint8)
    codes = np.zeros(len(out) * 4, dtype=np.uint8)
    codes[:len(a)] = (a >> 1) & 3
    q = codes.reshape(-1, 4)
    out[:] = q[:, 0] | q[:, 1] << 2 | q[:, 2] << 4 | q[:, 3] << 6
    return out


@pytest.mark.parametrize("scalar", [0, 1])
def test_pack2bit_host(lib, scalar):
    """The host half of the 2-bit transport form: AVX2 and table loop agree with the numpy restatement at every
    length around the 32-base groups and the 4 KiB validity blocks, from unaligned starts, and any byte outside
    ACGT -- wherever it sits -- refuses the range (it then travels raw)."""
    rng = np.random.default_rng(0xACE)
    src = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, 70000)].copy()
    for n in list(range(0, 70)) + [127, 128, 129, 4095, 4096, 4097, 4128, 8191, 65536, 69999]:
        for start in (0, 1, 13):
            a = src[start:start + n]
            dst = np.full(((len(a) + 15) // 16) * 4 + 8, 0xEE, dtype=np.uint8)
            assert lib.mgl_sw_pack2bit_host(a.ctypes.data, len(a), dst.ctypes.data, len(dst) - 8, scalar) == 1
            assert np.array_equal(dst[:-8], _pack2_numpy(a)), (n, start)
            assert (dst[-8:] == 0xEE).all()                    # nothing written past the packed words
    # unpacking with the kernel's table gives the bytes back
    a = src[:1000]
    packed = _pack2_numpy(a)
    codes = (packed[:, None] >> np.array([0, 2, 4, 6], dtype=np.uint8)) & 3
    assert np.array_equal(np.frombuffer(b"ACTG", dtype=np.uint8)[codes.reshape(-1)[:1000]], a)
    dst = np.zeros(70000, dtype=np.uint8)
    for pos in (0, 31, 32, 4095, 4096, 40000, 69998, 69999):
        for bad in (ord("N"), ord("a"), 0, 0x45, 0xC1):
            b = src[:70000].copy()
            b[pos] = bad
            assert lib.mgl_sw_pack2bit_host(b.ctypes.data, len(b), dst.ctypes.data, len(dst), scalar) == 0, (pos, bad)
    assert lib.mgl_sw_pack2bit_host(src.ctypes.data, 64, dst.ctypes.data, 15, scalar) == -1
    assert lib.mgl_sw_pack2bit_host(None, 64, dst.ctypes.data, 64, scalar) == -1


def test_plan_summary_buckets_config2(lib):
    from mgl_b200 import native, synth
    d = synth.generate(2, 4096)
    b, keep = native.make_batch(d["arena"], d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"])
    out = np.zeros(8, dtype=np.uint64)
    assert lib.mgl_sw_plan_summary(C.byref(b), 1, out.ctypes.data) == 0
    launches, units, n16, n32, trace_words, cells = (int(x) for x in out[:6])
    assert n16 == 4096 and n32 == 0                  # config 2 is 16-bit safe: two pairs per register
    assert units <= 4096 // 2 + 4 * launches + 64    # almost every unit holds two pairs
    assert cells == d["cells"]
    assert trace_words * 4 < 1.05 * cells * 1.2      # checkpoints: ~1.05 B/cell at K = 20 (sw_layout.h) plus padding
    assert lib.mgl_sw_plan_summary(C.byref(b), 0, out.ctypes.data) == 0
    assert int(out[2]) == 0 and int(out[3]) == 4096


def test_synth_is_a_pure_function_of_the_index():
    from mgl_b200 import synth
    a = synth.generate(5, 300, first=1000)
    b = synth.generate(5, 100, first=1100)
    for p in range(100):
        assert synth.pair_bytes(a, 100 + p) == synth.pair_bytes(b, p)
    c3 = synth.generate(3, 50)
    assert 300 <= c3["ref_len"].min() and c3["ref_len"].max() <= 1000


def test_no_cpu_fallback_without_a_gpu(lib):
    if lib.mgl_sw_device_count() > 0:
        pytest.skip("a GPU is present")
    import mgl_b200
    from mgl_b200 import native
    sw = mgl_b200.MicrosoftSmithWaterman()
    assert sw.load() is False                        # mirrors MicrosoftSmithWaterman.load: false, never throws
    with pytest.raises(native.MglSwError):
        sw.align(b"ACGT", b"ACGT", (10, -15, -30, -5), 1)
    with pytest.raises(native.MglSwError):
        mgl_b200.align_one_text(b"ACGT", b"ACGT", (10, -15, -30, -5), 1)


def test_pairhmm_library_exports_every_declared_symbol():
    import __graft_entry__ as g
    from mgl_b200 import pairhmm
    if not os.path.exists(pairhmm.LIB_PATH):
        g.build()
    L = pairhmm.lib()
    hdr = open(os.path.join(ROOT, "include", "mgl_pairhmm.h")).read()
    names = re.findall(r"MGL_PH_API\s+[\w\s\*]+?\b(mgl_pairhmm_\w+)\s*\(", hdr)
    assert len(names) >= 7
    for n in names:
        assert hasattr(L, n), n
    for n in ("initNative", "computeLikelihoodsNative", "doneNative"):     # ...MicrosoftPairHmm.h:15,23,31
        assert hasattr(L, "Java_com_microsoft_mgl_pairhmm_MicrosoftPairHmm_" + n)
    h = C.c_void_p()
    import ctypes
    if ctypes.CDLL(None) and L.mgl_pairhmm_create(0, C.byref(h)) != 0:
        assert pairhmm.MicrosoftPairHmm().load() is False        # no GPU: no fallback, load() reports it

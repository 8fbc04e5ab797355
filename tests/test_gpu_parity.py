"""GPU (B200): the CUDA path, called through the C-ABI, against the reference-
generated golden vectors, the oracle on seeded fuzz and synthetic configs, and
size-independent properties at BASELINE.json's full config-2 size.  Bit-exact:
identical CIGAR text and offset on every pair."""
import ctypes as C
import random
from collections import defaultdict

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def sw():
    import mgl_b200
    s = mgl_b200.MicrosoftSmithWaterman(0)
    s._ensure()                      # raises when libmgl_sw.so or the B200 is missing: nothing to fall back to
    yield s
    s.close()


def _check(sw, oracle, pairs, params, strategies, **opts):
    st = strategies if not isinstance(strategies, int) else None
    res = sw.align_batch(pairs, params, strategies if isinstance(strategies, int) else 1, st)
    bad = []
    for p, (r, a) in enumerate(pairs):
        s = strategies if isinstance(strategies, int) else int(strategies[p])
        exp = oracle.align(r, a, params, s)
        got = (res.cigar(p), int(res.offsets[p]))
        if got != exp:
            bad.append((p, len(r), len(a), params, s, got, exp))
    assert not bad, bad[:3]
    return res


@pytest.mark.parametrize("allow_s16", [1, 0])
def test_golden_vectors(sw, golden, allow_s16):
    sw.set_option("allow_s16", allow_s16)
    by = defaultdict(list)
    for c in golden["cases"]:
        by[tuple(c["params"])].append(c)
    total = 0
    try:
        for p, cs in by.items():
            pairs = [(c["ref"].encode("latin1"), c["alt"].encode("latin1")) for c in cs]
            for si, s in enumerate(golden["strategies"]):
                res = sw.align_batch(pairs, p, s)
                assert (res.stats["pairs_s16"] > 0) == bool(allow_s16)
                bad = [(c["ref"], c["alt"], p, s) for k, c in enumerate(cs)
                       if (res.cigar(k), int(res.offsets[k])) != tuple(c["expect"][si])]
                assert not bad, bad[:3]
                total += len(cs)
    finally:
        sw.set_option("allow_s16", 1)
    assert total > 10000


def test_single_pair_entry_points(sw, oracle):
    """the reference's one-pair call: Python mirror of MicrosoftSmithWaterman.align, mgl_sw_align_one, and the JNI symbol"""
    import mgl_b200
    from mgl_b200 import native
    P = mgl_b200.SWParameters(200, -150, -260, -11)
    r = sw.align(b"GGGGACGTACGTCCCC", b"TTACGTACGTTT", P, mgl_b200.SWOverhangStrategy.SOFTCLIP)
    assert (r.cigar, r.alignment_offset) == ("2I7M2I1M", 4)
    one = mgl_b200.align_one_text(b"ACGTACGTAC", b"GGGGGGGGGGGG", P, mgl_b200.SWOverhangStrategy.IGNORE)
    assert (one.cigar, one.alignment_offset) == ("13D1M", -11)

    # Java_..._alignNative with a table-backed JNIEnv whose GetDirectBufferAddress (slot 230) is the identity
    GETADDR = C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_void_p)
    ident = GETADDR(lambda env, buf: buf)
    table = (C.c_void_p * 235)()
    table[230] = C.cast(ident, C.c_void_p)
    env = C.pointer(C.c_void_p(C.addressof(table)))
    fn = native.lib().Java_com_microsoft_mgl_smithwaterman_MicrosoftSmithWaterman_alignNative
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_int] * 7
    fn.restype = C.c_int
    rng = random.Random(5)
    for k in range(40):
        ref = bytes(rng.choice(b"ACGT") for _ in range(rng.randint(1, 90)))
        alt = bytes(rng.choice(b"ACGT") for _ in range(rng.randint(1, 90)))
        s = (1, 2, 4, 8)[k % 4]
        reads = C.create_string_buffer(ref + alt, len(ref) + len(alt))
        cig = C.create_string_buffer(2 * max(len(ref), len(alt)) + 64)          # zero-filled like ByteBuffer.allocateDirect
        off = fn(env, None, C.addressof(reads), C.addressof(cig), len(ref), len(alt), 10, -15, -30, -5, s)
        assert (cig.value.decode().strip(), off) == oracle.align(ref, alt, (10, -15, -30, -5), s)


def test_fuzz_vs_oracle(sw, oracle):
    rng = random.Random(20261104)
    score_sets = [(10, -15, -30, -5), (200, -150, -260, -11), (1, -1, -1, -1), (3, -1, -2, 0), (2, -3, -1, -4)]
    for params in score_sets:
        pairs = []
        for k in range(1200):
            alpha = rng.choice([b"A", b"AC", b"ACGT", b"ACGTN", b"acgtACGT"])
            ref = bytes(rng.choice(alpha) for _ in range(rng.randint(1, 260)))
            if rng.random() < 0.5:
                i = rng.randrange(len(ref)); j = rng.randint(i + 1, len(ref))
                alt = bytearray(ref[i:j])
                for _ in range(rng.randint(0, 3)):
                    alt[rng.randrange(len(alt))] = rng.choice(alpha)
                alt = bytes(alt)
            else:
                alt = bytes(rng.choice(alpha) for _ in range(rng.randint(1, 260)))
            pairs.append((ref, alt))
        strategies = np.array([(1, 2, 4, 8)[k % 4] for k in range(len(pairs))], dtype=np.int32)
        _check(sw, oracle, pairs, params, strategies)


@pytest.mark.parametrize("params", [(10, -15, -30, -5), (200, -150, -260, -11), (2, -3, -1, -4)])
def test_fuzz_large_batches_through_the_round_kernels(sw, oracle, params):
    """batches of more than MGLSW_TRACE_ROUNDS_MIN_UNITS units: the short pairs' walks go from launch to launch
    (sw_trace_round_kernel, Walker::save / resume) -- all four strategies, degenerate alphabets, gaps at the edges,
    in 16-bit and in 32-bit arithmetic, in one chunk and in several"""
    rng = random.Random(hash(params) & 0xffff)
    pairs = []
    for k in range(7000):
        alpha = rng.choice([b"A", b"AC", b"ACGT", b"ACGTN"])
        ref = bytes(rng.choice(alpha) for _ in range(rng.randint(1, 220)))
        kind = rng.random()
        if kind < 0.6:
            i = rng.randrange(len(ref)); j = rng.randint(i + 1, len(ref))
            alt = bytearray(ref[i:j][:190])
            for _ in range(rng.randint(0, 4)):
                q = rng.randrange(len(alt))
                what = rng.random()
                if what < 0.4: alt[q] = rng.choice(alpha)
                elif what < 0.7 and len(alt) < 190: alt[q:q] = bytes(rng.choice(alpha) for _ in range(rng.randint(1, 9)))
                elif len(alt) > 12: del alt[q:q + rng.randint(1, 9)]
            alt = bytes(alt) or b"A"
        else:
            alt = bytes(rng.choice(alpha) for _ in range(rng.randint(1, 190)))
        pairs.append((ref, alt))
    strategies = np.array([(1, 2, 4, 8)[k % 4] for k in range(len(pairs))], dtype=np.int32)
    for chunk in (1 << 16, 3000):
        sw.set_option("chunk_pairs", chunk)
        try:
            res = _check(sw, oracle, pairs, params, strategies)
        finally:
            sw.set_option("chunk_pairs", 1 << 16)
        # the start kernel, five round launches and the final one on top of scan, scan, compact: per chunk (the last,
        # partial chunk may be small enough for the single launch)
        assert res.stats["other_launches"] >= 10 * max(1, res.stats["chunks"] - 1), res.stats


def test_long_gaps_take_the_gap_hops(sw, oracle):
    """deletions and insertions of 7 .. 200 cells at every kind of offset against the checkpoint grid (tests/gap_cases.py):
    the walk crosses whole blocks of such a gap with two checkpoint loads (Replay::ehop / fhop); as one small batch (one
    traceback launch, lane groups of 8 and whole-warp units side by side) and repeated into a batch large enough for
    the round kernels"""
    from gap_cases import long_gap_pairs
    pairs = long_gap_pairs()
    for params in ((10, -15, -30, -5), (200, -150, -260, -11), (3, -1, -2, 0)):
        strategies = np.array([(1, 2, 4, 8)[k % 4] for k in range(len(pairs))], dtype=np.int32)
        _check(sw, oracle, pairs, params, strategies)
        big = pairs * 8
        res = _check(sw, oracle, big, params, np.array([(2, 8, 1, 4)[k % 4] for k in range(len(big))], dtype=np.int32))
        assert res.stats["other_launches"] >= 10


def test_empty_and_invalid_inputs(sw):
    from mgl_b200 import native
    res = sw.align_batch([], (10, -15, -30, -5), 1)
    assert len(res.offsets) == 0 and int(res.cigar_start[0]) == 0
    with pytest.raises(native.MglSwError) as e:
        sw.align_batch([(b"ACGT", b"")], (10, -15, -30, -5), 1)
    assert e.value.code == native.E_PAIR
    res = sw.align_batch([(b"ACGT" * 50, b"ACGT" * 50)] * 3, (10, -15, -30, -5), 1)   # still usable after an error
    assert res.cigars() == ["200M"] * 3


def test_tiny_ops_buffer_reports_needed_size(sw):
    from mgl_b200 import native, synth
    d = synth.generate(2, 2000)
    b, keep = native.make_batch(d["arena"], d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, None)
    off, cs, ops = np.zeros(2000, np.int32), np.zeros(2001, np.uint64), np.zeros(10, np.uint32)
    r = native.Result()
    r.offsets, r.cigar_start, r.ops, r.ops_cap = off.ctypes.data, cs.ctypes.data, ops.ctypes.data, 10
    rc = native.lib().mgl_sw_align_batch(sw._ensure(), C.byref(b), C.byref(r), None)
    assert rc == native.E_OPS_CAPACITY and r.ops_total >= 2000 and int(cs[-1]) == r.ops_total


@pytest.mark.parametrize("config,n", [(2, 20000), (3, 3000), (5, 6000)])
def test_synthetic_configs_vs_oracle(sw, oracle, config, n):
    """the C restatement (oracle/sw_oracle.c) as the checker; pairs too large for its explicit matrices go to the
    reference's own align_scalar (oracle/_ref) -- none is skipped when that library is present"""
    from mgl_b200 import synth
    d = synth.generate(config, n)
    res = sw.align_arrays(d["arena"], d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"],
                          ops_cap=max(8 * n, int(d["ref_len"].sum() + d["alt_len"].sum()) // 8))
    assert res.stats["pairs"] == n
    step = 1 if n <= 3000 else 5
    big = np.nonzero(d["ref_len"].astype(np.int64) * d["alt_len"].astype(np.int64) > 4_000_000)[0]
    checked_big = 0
    for p in sorted(set(range(0, n, step)) | set(int(x) for x in big)):
        r, a = synth.pair_bytes(d, p)
        if len(r) * len(a) > 4_000_000:
            if not oracle.have_ref():
                continue
            exp = oracle.ref_align(r, a, d["params"], int(d["strategies"][p]), oracle.REF_SCALAR)
            checked_big += 1
        else:
            exp = oracle.align(r, a, d["params"], int(d["strategies"][p]))
        assert (res.cigar(p), int(res.offsets[p])) == exp, (config, p, len(r), len(a))
    if config == 5 and oracle.have_ref():
        assert checked_big == len(big) > 0            # every config-4-shaped pair of the mixed chunk was compared


# BASELINE.json configs at full size (config 4: 1 500 of its 10 000 pairs = 86 Gcells, config 5: 200 000 pairs as the
# stand-in for 100 M -- every long pair among them is compared), every pair against the reference's own C++
FULL_SIZE = [(2, 1_000_000), (3, 100_000), (4, 1_500), (5, 200_000)]


@pytest.mark.parametrize("config,n", FULL_SIZE)
def test_full_size_parity_vs_reference(sw, oracle, config, n):
    """SURVEY.md 8d 'parity at scale': the CUDA path through mgl_sw_align_batch (host buffers in and out) against
    oracle/_ref -- the reference's unmodified align_avx / align_scalar behind its JNI dispatch rule
    (...SmithWaterman.cpp:62-70) -- on EVERY pair, CIGAR text and offset bit-exact; then the same batch again over
    the 2-bit transport (pack_arena = 2), which must return identical arrays."""
    from mgl_b200 import synth
    if not oracle.have_ref():
        pytest.skip("oracle/_ref/libmgl_sw_ref.so not built (it is compiled in the authoring container and travels to the GPU box)")
    d = synth.generate(config, n)
    args = (d["arena"], d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"])
    cap = max(8 * n, int(d["ref_len"].astype(np.int64).sum() + d["alt_len"].astype(np.int64).sum()) // 8)
    res = sw.align_arrays(*args, ops_cap=cap)
    assert res.stats["pairs"] == n and res.stats["cells"] == d["cells"]
    if config == 5:
        long_pairs = int((d["alt_len"] > 4 * 32 * 24).sum())
        assert long_pairs >= 100 and res.stats["pairs_s16"] > 0 and res.stats["pairs_s32"] > long_pairs   # a really mixed batch
    nbad, first, secs = oracle.ref_check_batch(d["arena"], d["ref_off"], d["alt_off"], d["ref_len"], d["alt_len"], d["params"],
                                               d["strategies"], res.offsets, res.cigar_start, res.ops)
    detail = [(p, int(d["ref_len"][p]), int(d["alt_len"][p]), res.cigar(p)[:80], int(res.offsets[p]),
               oracle.ref_align(*synth.pair_bytes(d, p), d["params"], int(d["strategies"][p]))) for p in first[:3]]
    assert nbad == 0, (config, nbad, detail)
    sw.set_option("pack_arena", 2)
    try:
        packed = sw.align_arrays(*args, ops_cap=cap)
    finally:
        sw.set_option("pack_arena", 0)
    assert packed.stats["h2d_bytes"] < res.stats["h2d_bytes"]
    assert np.array_equal(packed.offsets, res.offsets) and np.array_equal(packed.cigar_start, res.cigar_start)
    assert np.array_equal(packed.ops, res.ops)


def test_packed_and_raw_transport_give_the_same_results(sw, oracle):
    """All-ACGT byte ranges may cross the host link at 2 bits per base and are expanded on the device (sw_pack.h); a
    range holding any other byte travels raw.  Same CIGARs and offsets either way, and equal to the oracle's: every
    chunk packed, none, chunks planned ahead of the upload only, and a batch with N / lower-case bytes in
    some chunks only (the reference compares raw bytes, SW/sw.cpp:55: N == N, a != A)."""
    from mgl_b200 import synth
    n = 24000
    d = synth.generate(2, n)
    args = (d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"])
    sw.set_option("chunk_pairs", 4096)                  # several chunks, each with its own byte range
    sw.set_option("host_plan", 1)                       # the 2-bit transport packs on the host planner's threads: compare like with like
    try:
        sw.set_option("pack_arena", 2)
        packed = sw.align_arrays(d["arena"], *args)
        sw.set_option("pack_arena", 0)
        raw = sw.align_arrays(d["arena"], *args)
        sw.set_option("pack_arena", 1)
        auto = sw.align_arrays(d["arena"], *args)
        sw.set_option("pack_arena", 2)
        assert packed.stats["h2d_bytes"] < 0.5 * raw.stats["h2d_bytes"]
        assert packed.stats["h2d_bytes"] <= auto.stats["h2d_bytes"] <= raw.stats["h2d_bytes"]
        assert packed.stats["other_launches"] > raw.stats["other_launches"]        # the unpack launches
        for r in (packed, auto):
            assert np.array_equal(r.offsets, raw.offsets) and np.array_equal(r.cigar_start, raw.cigar_start)
            assert np.array_equal(r.ops[:int(r.cigar_start[-1])], raw.ops[:int(raw.cigar_start[-1])])
        for p in range(0, n, 7):
            r, a = synth.pair_bytes(d, p)
            assert (packed.cigar(p), int(packed.offsets[p])) == oracle.align(r, a, d["params"], int(d["strategies"][p])), p
        # other bytes in the first, a middle and the last chunk: those chunks go up raw, the rest packed
        arena = d["arena"].copy()
        touched = []
        for p, byte in ((5, ord("N")), (6, ord("N")), (9000, ord("a")), (n - 1, ord("N")), (n - 2, ord("c"))):
            arena[int(d["ref_off"][p]) + 40 + len(touched)] = byte
            arena[int(d["alt_off"][p]) + 20 + len(touched)] = byte
            touched.append(p)
        mixed = sw.align_arrays(arena, *args)
        assert packed.stats["h2d_bytes"] < mixed.stats["h2d_bytes"] < 0.9 * raw.stats["h2d_bytes"]
        dm = dict(d, arena=arena)
        for p in sorted(set(touched + list(range(0, n, 11)))):
            r, a = synth.pair_bytes(dm, p)
            assert (mixed.cigar(p), int(mixed.offsets[p])) == oracle.align(r, a, d["params"], int(d["strategies"][p])), p
    finally:
        sw.set_option("pack_arena", 0)                  # the library defaults
        sw.set_option("host_plan", 0)
        sw.set_option("chunk_pairs", 1 << 16)


def test_long_pairs_strip_mined_vs_reference_or_oracle(sw, oracle):
    """config 4 shape (5-10 kbp): the whole-warp strip-mined kernel with HBM carry columns."""
    from mgl_b200 import synth
    d = synth.generate(4, 6)
    res = sw.align_arrays(d["arena"], d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"])
    assert res.stats["pairs_s32"] == 6
    for p in range(6):
        r, a = synth.pair_bytes(d, p)
        if oracle.have_ref():
            exp = oracle.ref_align(r, a, d["params"], 1, oracle.REF_SCALAR)      # the reference's own align_scalar
        elif len(r) * len(a) <= 40_000_000:
            exp = oracle.align(r, a, d["params"], 1)
        else:
            continue
        assert (res.cigar(p), int(res.offsets[p])) == exp, (p, len(r), len(a))


def test_full_size_config2_properties(sw):
    """1 M pairs (BASELINE.json config 2): both arithmetics agree on every pair; CIGARs consume exactly the alt and
    stay inside the ref; resident (kernel-timing) path returns the same results as the host-buffer path."""
    from mgl_b200 import synth
    n = 1_000_000
    d = synth.generate(2, n)
    args = (d["arena"], d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"])
    a = sw.align_arrays(*args)
    assert a.stats["pairs_s16"] == n
    sw.set_option("allow_s16", 0)
    try:
        b = sw.align_arrays(*args)
    finally:
        sw.set_option("allow_s16", 1)
    assert b.stats["pairs_s32"] == n
    assert np.array_equal(a.offsets, b.offsets) and np.array_equal(a.cigar_start, b.cigar_start) and np.array_equal(a.ops, b.ops)
    sw.resident_prepare(*args)
    st = sw.resident_run()
    assert st["cells"] == d["cells"]
    c = sw.resident_fetch()
    sw.resident_drop()
    assert np.array_equal(a.offsets, c.offsets) and np.array_equal(a.ops, c.ops)
    # CIGAR arithmetic: M + I + S == alt length; offset + M + D <= ref length
    lens, op = (a.ops >> 4).astype(np.int64), a.ops & 15
    pair_of = np.repeat(np.arange(n), np.diff(a.cigar_start).astype(np.int64))
    q = np.bincount(pair_of, weights=lens * np.isin(op, (0, 1, 4)), minlength=n)
    t = np.bincount(pair_of, weights=lens * np.isin(op, (0, 2)), minlength=n)
    assert np.array_equal(q.astype(np.int64), d["alt_len"].astype(np.int64))
    assert (a.offsets >= 0).all() and (a.offsets + t <= d["ref_len"]).all()


def test_concurrent_single_pair_calls_are_coalesced(sw, oracle):
    """the reference's entry is one pair per call from many JVM threads (static native, re-entrant): concurrent
    mgl_sw_align_one calls must all return the reference's answer (they are coalesced into batches underneath)"""
    import threading
    import mgl_b200
    rng = random.Random(77)
    jobs = []
    for k in range(16 * 60):
        ref = bytes(rng.choice(b"ACGT") for _ in range(rng.randint(20, 300)))
        i = rng.randrange(len(ref)); j = rng.randint(i + 1, len(ref))
        alt = bytearray(ref[i:j]) + bytes(rng.choice(b"ACGT") for _ in range(rng.randint(0, 6)))
        if len(alt) > 3:
            alt[rng.randrange(len(alt))] = rng.choice(b"ACGT")
        params = ((10, -15, -30, -5), (200, -150, -260, -11))[k % 2]
        jobs.append((ref, bytes(alt), params, (1, 2, 4, 8)[k % 4]))
    got = [None] * len(jobs)

    def work(t):
        for k in range(t, len(jobs), 16):
            r, a, p, s = jobs[k]
            one = mgl_b200.align_one_text(r, a, p, s)
            got[k] = (one.cigar, one.alignment_offset)

    threads = [threading.Thread(target=work, args=(t,)) for t in range(16)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for k, (r, a, p, s) in enumerate(jobs):
        assert got[k] == oracle.align(r, a, p, s), (k, len(r), len(a), p, s)


def test_pipelined_strips_edge_shapes(sw, oracle):
    """pairs of >= 4 strips run one per block with the strips pipelined over its warps (sw_fill_pipe.cuh): refs shorter
    than the lane skew, the 3 / 4 strip boundary (alt 2304 / 2305), every strategy"""
    rng = random.Random(4242)
    pairs = []
    for tl, ql in ((1, 3000), (5, 2500), (31, 2305), (33, 2400), (40, 3100), (700, 2304), (650, 2305), (1500, 3073), (900, 4000)):
        ref = bytes(rng.choice(b"ACGT") for _ in range(tl))
        alt = bytearray(rng.choice(b"ACGT") for _ in range(ql))
        w = min(tl, ql)
        start = rng.randrange(0, ql - w + 1)
        alt[start:start + w] = ref[:w]
        for _ in range(8):
            alt[rng.randrange(ql)] = rng.choice(b"ACGT")
        pairs.append((ref, bytes(alt)))
    try:
        for wide in (0, 2):                       # four warps per pair (throughput shape), eight (latency shape)
            sw.set_option("pipe_wide", wide)
            for params in ((200, -150, -260, -11), (10, -15, -30, -5)):
                strategies = np.array([(1, 2, 4, 8)[k % 4] for k in range(len(pairs))], dtype=np.int32)
                res = _check(sw, oracle, pairs, params, strategies)
                strategies = np.array([(2, 8, 1, 4)[k % 4] for k in range(len(pairs))], dtype=np.int32)
                _check(sw, oracle, pairs, params, strategies)
            assert res.stats["pairs_s32"] >= 7
    finally:
        sw.set_option("pipe_wide", 1)


def test_score_only_batch(sw, oracle, golden):
    """mgl_sw_score_batch (no checkpoints, no traceback): (score, row, column) of the traceback's start cell against the
    oracle (pinned to the reference's ScoreMax in tests/test_oracle.py): golden cases, both arithmetics, lane groups of
    8 and 32, strip-mined and pipelined pairs, every strategy"""
    rng = random.Random(777)
    pairs = [(c["ref"].encode("latin1"), c["alt"].encode("latin1")) for c in golden["cases"][:400]]
    for tl, ql in ((1, 1), (7, 300), (300, 7), (450, 150), (500, 193), (900, 700), (640, 1000), (33, 2400), (800, 2305), (1200, 3100)):
        ref = bytes(rng.choice(b"ACGT") for _ in range(tl))
        alt = bytearray(rng.choice(b"ACGT") for _ in range(ql))
        w = min(tl, ql)
        start = rng.randrange(0, ql - w + 1)
        alt[start:start + w] = ref[:w]
        for _ in range(6):
            alt[rng.randrange(ql)] = rng.choice(b"ACGT")
        pairs.append((ref, bytes(alt)))
    for params in ((10, -15, -30, -5), (200, -150, -260, -11)):
        for rot in range(2):
            strategies = np.array([(1, 2, 4, 8)[(k + rot) % 4] for k in range(len(pairs))], dtype=np.int32)
            got = sw.score_batch(pairs, params, 1, strategies)
            bad = []
            for k, (r, a) in enumerate(pairs):
                exp = oracle.score3(r, a, params, int(strategies[k]))
                if tuple(int(x) for x in got[k]) != exp:
                    bad.append((k, len(r), len(a), params, int(strategies[k]), tuple(int(x) for x in got[k]), exp))
            assert not bad, bad[:3]
    # the same handle still aligns afterwards (slots switch back to checkpointed runs)
    _check(sw, oracle, pairs[-10:], (200, -150, -260, -11), np.array([1, 2, 4, 8, 1, 2, 4, 8, 1, 2], dtype=np.int32))


def test_score_only_matches_alignment_on_synthetic_config(sw):
    """config 2 and 3 shapes, 20 000 pairs: the score-only start cell agrees with the alignment's own end point
    (offset + ref bases consumed = start row; alt bases before the trailing clip = start column)"""
    from mgl_b200 import synth
    for config in (2, 3):
        d = synth.generate(config, 20000)
        res = sw.align_arrays(d["arena"], d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"],
                              ops_cap=16 * 20000)
        sc, _ = sw.score_arrays(d["arena"], d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"])
        cs = res.cigar_start.astype(np.int64)
        for p in range(0, 20000, 7):
            ops = res.ops[cs[p]: cs[p + 1]]
            ref_used = sum(int(x) >> 4 for x in ops if (int(x) & 15) in (0, 2))
            alt_used = sum(int(x) >> 4 for x in ops if (int(x) & 15) in (0, 1, 4))
            strat = int(d["strategies"][p])
            trail = int(ops[-1]) >> 4 if (int(ops[-1]) & 15) == 4 and strat == 1 and len(ops) > 1 else 0
            assert int(res.offsets[p]) + ref_used == int(sc[p, 1]), (config, p)
            assert alt_used - trail == int(sc[p, 2]), (config, p)


def test_text_cigars_formatted_on_device(sw, oracle, golden):
    """mgl_sw_align_batch_text: the characters alignNative writes (sw.cpp:251-252), formatted by the GPU, against the
    oracle's text -- golden cases, long elements (4- and 5-digit lengths), a capacity retry"""
    rng = random.Random(99)
    pairs = [(c["ref"].encode("latin1"), c["alt"].encode("latin1")) for c in golden["cases"][:600]]
    for tl, ql in ((1200, 1100), (12000, 25), (30, 11000), (2500, 2400)):
        ref = bytes(rng.choice(b"ACGT") for _ in range(tl))
        w = min(tl, ql)
        alt = bytearray(ref[:w]) + bytearray(rng.choice(b"ACGT") for _ in range(ql - w))
        pairs.append((ref, bytes(alt)))
    for params in ((10, -15, -30, -5), (200, -150, -260, -11)):
        strategies = np.array([(1, 2, 4, 8)[k % 4] for k in range(len(pairs))], dtype=np.int32)
        got = sw.align_batch_text(pairs, params, 1, strategies)
        bad = [(k, got[k], oracle.align(r, a, params, int(strategies[k]))) for k, (r, a) in enumerate(pairs)
               if got[k] != oracle.align(r, a, params, int(strategies[k]))]
        assert not bad, bad[:3]
    # capacity: too small a text buffer reports what is needed and the retry succeeds
    from mgl_b200.smithwaterman import _pack_pairs
    arena, ro, rl, ao, al = _pack_pairs(pairs[:50])
    offsets, start, text, _ = sw.align_arrays_text(arena, ro, rl, ao, al, (10, -15, -30, -5), 1, None, text_cap=8)
    assert int(start[-1]) == len(text) > 8


def test_jni_batch_symbol(sw, oracle):
    """Java_..._alignBatchNative through a table-backed JNIEnv (GetDirectBufferAddress = identity): what a Java
    alignBatch would receive, against the oracle"""
    from mgl_b200 import native
    from mgl_b200.smithwaterman import _pack_pairs
    GETADDR = C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_void_p)
    ident = GETADDR(lambda env, buf: buf)
    table = (C.c_void_p * 235)()
    table[230] = C.cast(ident, C.c_void_p)
    env = C.pointer(C.c_void_p(C.addressof(table)))
    fn = native.lib().Java_com_microsoft_mgl_smithwaterman_MicrosoftSmithWaterman_alignBatchNative
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int] + [C.c_void_p] * 4 + [C.c_int] * 5 + [C.c_void_p] * 4 + [C.c_int64]
    fn.restype = C.c_int
    rng = random.Random(31)
    pairs = []
    for _ in range(300):
        ref = bytes(rng.choice(b"ACGT") for _ in range(rng.randint(20, 400)))
        i = rng.randrange(len(ref) - 10)
        alt = bytearray(ref[i: i + rng.randint(10, 150)])
        for _ in range(rng.randint(0, 3)):
            alt[rng.randrange(len(alt))] = rng.choice(b"ACGT")
        pairs.append((ref, bytes(alt)))
    arena, ro, rl, ao, al = _pack_pairs(pairs)
    n = len(pairs)
    strategies = np.array([(1, 2, 4, 8)[k % 4] for k in range(n)], dtype=np.int32)
    offsets, start, text = np.zeros(n, np.int32), np.zeros(n + 1, np.uint64), np.zeros(64 * n, np.uint8)
    rc = fn(env, None, arena.ctypes.data, arena.nbytes, n, ro.ctypes.data, rl.ctypes.data, ao.ctypes.data, al.ctypes.data,
            10, -15, -30, -5, 1, strategies.ctypes.data, offsets.ctypes.data, start.ctypes.data, text.ctypes.data, text.nbytes)
    assert rc == 0
    raw = text.tobytes()
    for k, (r, a) in enumerate(pairs):
        assert (raw[int(start[k]): int(start[k + 1])].decode(), int(offsets[k])) == oracle.align(r, a, (10, -15, -30, -5), int(strategies[k])), k
    # too small a text buffer: the code and the needed capacity
    rc = fn(env, None, arena.ctypes.data, arena.nbytes, n, ro.ctypes.data, rl.ctypes.data, ao.ctypes.data, al.ctypes.data,
            10, -15, -30, -5, 1, None, offsets.ctypes.data, start.ctypes.data, text.ctypes.data, 16)
    assert rc == native.E_OPS_CAPACITY and int(start[n]) > 16


def test_multi_device_handle_matches_single_and_oracle(sw, oracle):
    """mgl_sw_create_multi: one handle over a device list cuts the batch by DP cells, runs one
    engine per entry and gathers in input order.  Output == the one-device handle's == the reference's on a 200 000-pair
    config-5 batch (short, medium and long pairs in every shard), for packed elements, device-formatted text and the
    score-only entry.  The list is every visible GPU, or device 0 three times on a one-GPU box (three engines sharing
    it: the sharding, staging and gather logic is the same)."""
    import mgl_b200
    from mgl_b200 import native, synth
    ndev = native.lib().mgl_sw_device_count()
    devices = list(range(ndev)) if ndev > 1 else [0, 0, 0]
    n = 200_000
    d = synth.generate(5, n)
    args = (d["arena"], d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"])
    cap = max(8 * n, int(d["ref_len"].astype(np.int64).sum() + d["alt_len"].astype(np.int64).sum()) // 8)
    single = sw.align_arrays(*args, ops_cap=cap)
    multi = mgl_b200.MicrosoftSmithWaterman(devices)
    try:
        multi._ensure()
        if ndev == 1:
            multi.set_option("trace_budget_mb", 3000)     # three engines (and the fixture's) share one GPU's memory here
        got = (C.c_int * 16)()
        assert native.lib().mgl_sw_handle_devices(multi._h, got, 16) == len(devices) and list(got[:len(devices)]) == devices
        res = multi.align_arrays(*args, ops_cap=cap)
        assert res.stats["pairs"] == n and res.stats["cells"] == d["cells"]
        assert np.array_equal(res.offsets, single.offsets) and np.array_equal(res.cigar_start, single.cigar_start)
        assert np.array_equal(res.ops, single.ops)
        if oracle.have_ref():
            nbad, first, _ = oracle.ref_check_batch(d["arena"], d["ref_off"], d["alt_off"], d["ref_len"], d["alt_len"], d["params"],
                                                    d["strategies"], res.offsets, res.cigar_start, res.ops)
            assert nbad == 0, (nbad, first)
        # too small an output buffer: the needed size comes back, as from one device
        small = np.zeros(1000, dtype=np.uint32)
        r = native.Result()
        off, cs = np.zeros(n, np.int32), np.zeros(n + 1, np.uint64)
        r.offsets, r.cigar_start, r.ops, r.ops_cap = off.ctypes.data, cs.ctypes.data, small.ctypes.data, len(small)
        b, keep = native.make_batch(*args[:5], d["params"], 1, d["strategies"])
        assert native.lib().mgl_sw_align_batch(multi._h, C.byref(b), C.byref(r), None) == native.E_OPS_CAPACITY
        assert r.ops_total == len(single.ops)
        # text and score-only through the same sharding
        m = 30_000
        sub = tuple(a[:m] if k in (1, 2, 3, 4) else a for k, a in enumerate(args[:5])) + (d["params"], 1, d["strategies"][:m])
        t1 = sw.align_arrays_text(*sub)
        t2 = multi.align_arrays_text(*sub)
        assert np.array_equal(t1[0], t2[0]) and np.array_equal(t1[1], t2[1]) and np.array_equal(t1[2], t2[2])
        s1, _ = sw.score_arrays(*sub)
        s2, _ = multi.score_arrays(*sub)
        assert np.array_equal(s1, s2)
        # edge: fewer pairs than devices, and an empty batch
        tiny = multi.align_batch([(b"ACGTACGTAC", b"ACGTTCGTAC")], (10, -15, -30, -5), 1)
        assert tiny.cigars() == ["10M"]
        assert len(multi.align_batch([], (10, -15, -30, -5), 1).offsets) == 0
    finally:
        multi.close()


def test_default_handle_survives_done_native(sw, oracle):
    """doneNative (MicrosoftSmithWaterman.close(), .java:89-92) releases the process-wide handle; the next alignNative
    creates it again"""
    import mgl_b200
    from mgl_b200 import native
    P = (200, -150, -260, -11)
    one = mgl_b200.align_one_text(b"GGGGACGTACGTCCCC", b"TTACGTACGTTT", P, 1)
    assert (one.cigar, one.alignment_offset) == ("2I7M2I1M", 4)
    fn = native.lib().Java_com_microsoft_mgl_smithwaterman_MicrosoftSmithWaterman_doneNative
    fn.argtypes, fn.restype = [C.c_void_p, C.c_void_p], None
    fn(None, None)
    fn(None, None)                                       # idempotent
    one = mgl_b200.align_one_text(b"ACGTACGTAC", b"GGGGGGGGGGGG", P, 8)
    assert (one.cigar, one.alignment_offset) == ("13D1M", -11)


def test_device_planner_and_host_planner_give_identical_results(sw, oracle):
    """Chunks are planned by device kernels (sw_plan_dev.cu) unless "host_plan" is set: the same batches through both
    planners must return identical arrays -- a mixed config-5 batch (every kernel shape, several chunks), tie-heavy
    fuzz with all four strategies, text and score-only entries -- and bad input must fail the same way."""
    from mgl_b200 import native, synth
    n = 120_000
    d = synth.generate(5, n)
    args = (d["arena"], d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"])
    cap = max(8 * n, int(d["ref_len"].astype(np.int64).sum() + d["alt_len"].astype(np.int64).sum()) // 8)
    rng = random.Random(4)
    fuzz = []
    for k in range(3000):
        alpha = rng.choice([b"A", b"AC", b"ACGT"])
        ref = bytes(rng.choice(alpha) for _ in range(rng.randint(1, 400)))
        alt = bytes(rng.choice(alpha) for _ in range(rng.randint(1, 260)))
        fuzz.append((ref, alt))
    fst = np.array([(1, 2, 4, 8)[k % 4] for k in range(len(fuzz))], dtype=np.int32)
    got = {}
    try:
        for hp in (0, 1):
            sw.set_option("host_plan", hp)
            sw.set_option("chunk_pairs", 16384)
            a = sw.align_arrays(*args, ops_cap=cap)
            b = sw.align_batch(fuzz, (3, -1, -2, 0), 1, fst)
            sw.set_option("allow_s16", 0)
            c = sw.align_batch(fuzz, (3, -1, -2, 0), 1, fst)
            sw.set_option("allow_s16", 1)
            t = sw.align_arrays_text(*(x[:20000] if k in (1, 2, 3, 4) else x for k, x in enumerate(args[:5])), d["params"], 1, d["strategies"][:20000])
            s3, _ = sw.score_arrays(*(x[:20000] if k in (1, 2, 3, 4) else x for k, x in enumerate(args[:5])), d["params"], 1, d["strategies"][:20000])
            got[hp] = (a, b, c, t, s3)
            assert a.stats["pairs"] == n and a.stats["cells"] == d["cells"] and a.stats["chunks"] > 4
            with pytest.raises(native.MglSwError) as e:
                sw.align_batch([(b"ACGT", b"ACGT")] * 5 + [(b"", b"ACGT")], (10, -15, -30, -5), 1)
            assert e.value.code == native.E_PAIR
            bad = native.make_batch(np.frombuffer(b"ACGTACGT", dtype=np.uint8), [0, 2 ** 63], [4, 4], [4, 4], [4, 4], (10, -15, -30, -5), 1, None)
            r = native.Result()
            off, cs, ops = np.zeros(2, np.int32), np.zeros(3, np.uint64), np.zeros(64, np.uint32)
            r.offsets, r.cigar_start, r.ops, r.ops_cap = off.ctypes.data, cs.ctypes.data, ops.ctypes.data, 64
            assert native.lib().mgl_sw_align_batch(sw._ensure(), C.byref(bad[0]), C.byref(r), None) == native.E_INVALID
    finally:
        sw.set_option("host_plan", 0)
        sw.set_option("chunk_pairs", 1 << 16)
        sw.set_option("allow_s16", 1)
    for x, y in zip(got[0][:3], got[1][:3]):
        assert np.array_equal(x.offsets, y.offsets) and np.array_equal(x.cigar_start, y.cigar_start) and np.array_equal(x.ops, y.ops)
        assert x.stats["pairs_s16"] == y.stats["pairs_s16"] and x.stats["fill_launches"] == y.stats["fill_launches"]
        assert x.stats["trace_bytes"] == y.stats["trace_bytes"]
    for k in range(3):
        assert np.array_equal(got[0][3][k], got[1][3][k])
    assert np.array_equal(got[0][4], got[1][4])
    for p in range(0, len(fuzz), 3):
        exp = oracle.align(fuzz[p][0], fuzz[p][1], (3, -1, -2, 0), int(fst[p]))
        assert (got[0][1].cigar(p), int(got[0][1].offsets[p])) == exp and (got[0][2].cigar(p), int(got[0][2].offsets[p])) == exp, p


def test_exact_match_shortcut_is_opt_in_and_matches_gatks_wrapper(sw, oracle):
    """SURVEY.md 8f row 2.  Default: the reference's own output for every pair.  With "exact_match_shortcut": SOFTCLIP and
    IGNORE pairs whose alt occurs verbatim in the ref return what GATK's SWNativeAlignerWrapper returns before it would
    call the binding (...MicrosoftSmithWaterman.cpp:60 leaves the check to it): offset = the LAST occurrence, CIGAR = all M;
    INDEL / LEADING_INDEL pairs and pairs without an exact occurrence are untouched."""
    rng = random.Random(606)
    pairs = []
    for k in range(600):
        unit = bytes(rng.choice(b"ACGT") for _ in range(rng.randint(3, 40)))
        ref = (unit * rng.randint(1, 8) + bytes(rng.choice(b"ACGT") for _ in range(rng.randint(0, 60))))[: rng.randint(20, 400)]
        if k % 3 == 0:
            i = rng.randrange(len(ref)); j = rng.randint(i + 1, min(len(ref), i + 200))
            alt = ref[i:j]                                            # verbatim, often repeated in the ref
        elif k % 3 == 1:
            alt = bytes(rng.choice(b"ACGT") for _ in range(rng.randint(1, 150)))
        else:
            i = rng.randrange(len(ref)); j = rng.randint(i + 1, min(len(ref), i + 200))
            a = bytearray(ref[i:j]); a[rng.randrange(len(a))] = ord("N"); alt = bytes(a)
        pairs.append((ref, alt))
    params = (10, -15, -30, -5)
    strategies = np.array([(1, 8, 2, 4)[k % 4] for k in range(len(pairs))], dtype=np.int32)
    plain = sw.align_batch(pairs, params, 1, strategies)
    sw.set_option("exact_match_shortcut", 1)
    try:
        short = sw.align_batch(pairs, params, 1, strategies)
        text = sw.align_batch_text(pairs, params, 1, strategies)
    finally:
        sw.set_option("exact_match_shortcut", 0)
    hits = differ = 0
    for k, (r, a) in enumerate(pairs):
        s = int(strategies[k])
        dp = oracle.align(r, a, params, s)
        assert (plain.cigar(k), int(plain.offsets[k])) == dp, k
        at = r.rfind(a) if s in (1, 8) else -1
        exp = (f"{len(a)}M", at) if at >= 0 else dp
        assert (short.cigar(k), int(short.offsets[k])) == exp, (k, s, len(r), len(a), at)
        assert text[k] == exp, k
        hits += at >= 0
        differ += at >= 0 and exp != dp
    assert hits > 60                           # the shortcut fired; on these cases the DP's own tie rules (last-column maximum
    del differ                                 # with ">=", sw.cpp:100-104) pick the last occurrence too: `differ` stays 0


def test_work_slots_follow_the_batch_kind(sw, oracle):
    """two work slots for long pairs, three for mixed batches, four for short pairs: one handle alternates between them
    (buffers of the other layout are released first) and every batch still matches the reference"""
    from mgl_b200 import synth
    for config, n in ((2, 30000), (4, 24), (5, 40000), (3, 4000), (2, 30000)):
        d = synth.generate(config, n)
        cap = max(8 * n, int(d["ref_len"].astype(np.int64).sum() + d["alt_len"].astype(np.int64).sum()) // 8)
        res = sw.align_arrays(d["arena"], d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"], ops_cap=cap)
        if oracle.have_ref():
            nbad, first, _ = oracle.ref_check_batch(d["arena"], d["ref_off"], d["alt_off"], d["ref_len"], d["alt_len"], d["params"],
                                                    d["strategies"], res.offsets, res.cigar_start, res.ops)
            assert nbad == 0, (config, nbad, first)
    for slots in (2, 3, 4):
        sw.set_option("slots", slots)
        d = synth.generate(3, 3000)
        res = sw.align_arrays(d["arena"], d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"], ops_cap=64 * 3000)
        for p in range(0, 3000, 41):
            r, a = synth.pair_bytes(d, p)
            assert (res.cigar(p), int(res.offsets[p])) == oracle.align(r, a, d["params"], int(d["strategies"][p])), (slots, p)
    sw.set_option("slots", 0)

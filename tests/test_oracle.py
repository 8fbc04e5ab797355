"""CPU: pin the oracle restatement (oracle/sw_oracle.c) against the reference.

1. against the committed golden vectors, which are outputs of the reference's
   own align_scalar/align_avx/JNI entry (tests/golden/make_golden.py);
2. when oracle/_ref exists (authoring container, or shipped prebuilt to the GPU
   box), differentially against the reference binary on fresh seeded fuzz.
"""
import random

import pytest


def test_oracle_matches_golden(golden, oracle):
    bad = []
    for c in golden["cases"]:
        r, a, p = c["ref"].encode("latin1"), c["alt"].encode("latin1"), tuple(c["params"])
        for s, (cig, off) in zip(golden["strategies"], c["expect"]):
            got = oracle.align(r, a, p, s)
            if got != (cig, off):
                bad.append((c["ref"], c["alt"], p, s, got, (cig, off)))
    assert not bad, bad[:5]
    assert len(golden["cases"]) > 2000


def test_appendix_b_known_answers(oracle):
    # SURVEY.md Appendix B, as offset:CIGAR for SOFTCLIP, INDEL, LEADING_INDEL, IGNORE
    P1, P2 = (10, -15, -30, -5), (200, -150, -260, -11)
    table = [
        (b"ACGTTTTTTGCA", b"ACGTTTGCAGGG", P1, ["0:3M3D6M3S", "0:3M3D6M3I", "0:3M3D6M3I", "0:3M3D9M"]),
        (b"GGGGACGTACGTCCCC", b"TTACGTACGTTT", P2, ["4:2I7M2I1M", "0:4D2I8M4D2I", "0:4D2I7M2I1M", "4:2I7M2I1M"]),
        (b"ACGTACGTAC", b"GGGGGGGGGGGG", P2, ["0:11S2D1M", "0:10D12I", "0:2D11I1M", "-11:13D1M"]),
        (b"ACTGACTGACTGAAAAACTG", b"ACTGACTGACTGACTG", P1, ["0:4S12M", "0:12M4D4M", "0:12M4D4M", "-4:16M"]),
    ]
    for r, a, p, exp in table:
        got = ["%d:%s" % (oracle.align(r, a, p, s)[1], oracle.align(r, a, p, s)[0]) for s in oracle.STRATEGIES]
        assert got == exp


def test_oracle_vs_reference_binary(oracle):
    if not oracle.have_ref():
        pytest.skip("oracle/_ref not built (no /root/reference here); golden vectors cover this")
    rng = random.Random(20261104)
    score_sets = [(10, -15, -30, -5), (200, -150, -260, -11), (1, -1, -1, -1), (3, -1, -2, 0), (2, -3, -1, -4)]
    n = 0
    for k in range(1500):
        alpha = rng.choice(["A", "AC", "ACGT", "ACGTN"])
        ref = bytes(rng.choice(alpha.encode()) for _ in range(rng.randint(1, 120)))
        if rng.random() < 0.5:
            i = rng.randrange(len(ref)); j = rng.randint(i + 1, len(ref))
            alt = bytearray(ref[i:j])
            for _ in range(rng.randint(0, 3)):
                alt[rng.randrange(len(alt))] = rng.choice(alpha.encode())
            alt = bytes(alt)
        else:
            alt = bytes(rng.choice(alpha.encode()) for _ in range(rng.randint(1, 120)))
        p = score_sets[k % len(score_sets)]
        for s in oracle.STRATEGIES:
            assert oracle.align(ref, alt, p, s) == oracle.ref_align(ref, alt, p, s), (ref, alt, p, s)
            n += 1
    assert n == 6000


def test_reference_jni_entry_matches_oracle(oracle):
    if not oracle.have_ref():
        pytest.skip("oracle/_ref not built")
    r, a = b"GGGGACGTACGTCCCC", b"TTACGTACGTTT"
    assert oracle.ref_align_jni(r, a, (200, -150, -260, -11), oracle.SOFTCLIP) == ("2I7M2I1M", 4)
    assert oracle.align(r, a, (200, -150, -260, -11), oracle.SOFTCLIP) == ("2I7M2I1M", 4)


def test_start_cell_and_score_vs_reference_scoremax(oracle):
    """the oracle's (score, row, column) of the traceback start cell -- what mgl_sw_score_batch returns -- against the
    reference's own ScoreMax (its calculateMatrix, sw.cpp:5-146) and against the score of the reference's output CIGAR"""
    if not oracle.have_ref():
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    rng = random.Random(8104)
    score_sets = [(10, -15, -30, -5), (200, -150, -260, -11), (1, -1, -1, -1), (3, -1, -2, 0)]   # open >= extend: CIGAR runs = gaps
    for k in range(600):
        alpha = rng.choice(["AC", "ACGT", "ACGTN"])
        ref = bytes(rng.choice(alpha.encode()) for _ in range(rng.randint(1, 150)))
        if rng.random() < 0.6:
            i = rng.randrange(len(ref)); j = rng.randint(i + 1, len(ref))
            alt = bytearray(ref[i:j])
            for _ in range(rng.randint(0, 4)):
                alt[rng.randrange(len(alt))] = rng.choice(alpha.encode())
            alt = bytes(alt) + bytes(rng.choice(alpha.encode()) for _ in range(rng.randint(0, 6)))
        else:
            alt = bytes(rng.choice(alpha.encode()) for _ in range(rng.randint(1, 150)))
        p = score_sets[k % len(score_sets)]
        for s in oracle.STRATEGIES:
            score, si, sj = oracle.score3(ref, alt, p, s)
            ez = oracle.ref_scoremax(ref, alt, p, s)
            if s == oracle.INDEL:
                assert (si, sj) == (len(ref), len(alt))
            elif s == oracle.LEADING_INDEL:
                assert (score, si, sj) == (ez["mqe"], ez["mqe_t"], len(alt)), (ref, alt, p, s)
            else:
                assert (score, si, sj) == (ez["max"], ez["max_t"], ez["max_q"]), (ref, alt, p, s)
            if s != oracle.IGNORE:          # IGNORE folds the overhangs into M: its CIGAR is not a scored path
                cigar, off = oracle.ref_align(ref, alt, p, s)
                assert oracle.path_score(cigar, off, ref, alt, p) == score, (ref, alt, p, s, cigar, off)

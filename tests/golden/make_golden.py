#!/usr/bin/env python
"""Generate tests/golden/sw_golden.json.gz from the REFERENCE ITSELF.

Runs in the authoring container only (needs /root/reference to have been
compiled into oracle/_ref by `make -C oracle ref`).  Every expected (cigar,
offset) below is the output of the reference's own align_scalar (sw.cpp:258),
cross-checked against its align_avx (sw_avx.cpp:6) and its JNI entry point.
The GPU box has no /root/reference; tests read only the committed fixture.

Cases:
  kat      SURVEY.md Appendix B sequences, 2 score sets x 4 strategies
  pairhmm  hap/read columns of the reference's src/test/resources/pairhmm-testdata.txt
           (104 realistic pairs; the file carries no SW answers), 2 x 4
  fuzz     seeded random pairs: 1-4 letter alphabets (tie heavy), exact substrings,
           mutated copies, non-ACGT bytes, 7 score sets incl. gap_extend = 0 and
           gap_open < gap_extend, all 4 strategies
"""
import gzip
import json
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import sw_oracle as o  # noqa: E402

P_READ = (10, -15, -30, -5)        # BASELINE.json config 2
P_HAP = (200, -150, -260, -11)     # BASELINE.json config 3/4
SCORE_SETS = [P_READ, P_HAP, (1, -1, -1, -1), (3, -1, -2, 0), (5, -4, -1, -3), (2, -3, -5, -2), (25, -50, -110, -6)]

KATS = [("ACGTACGTACGT", "ACGTACGT"), ("AAAAAAAAAAAA", "AAAAAAAA"), ("ACGTTTTTTGCA", "ACGTTTGCAGGG"),
        ("GGGGACGTACGTCCCC", "TTACGTACGTTT"), ("ACGTACGTAC", "GGGGGGGGGGGG"),
        ("ACTGACTGACTGAAAAACTG", "ACTGACTGACTGACTG"), ("A", "A"), ("A", "C"), ("ACGT", "A"), ("A", "ACGT"),
        ("AC", "CA"), ("NNNNACGTNNNN", "nnnnACGTnnnn")]


def ref_all(r: bytes, a: bytes, p, s):
    sc = o.ref_align(r, a, p, s, o.REF_SCALAR)
    jn = o.ref_align_jni(r, a, p, s)
    assert sc == jn, (r, a, p, s, sc, jn)
    if len(a) >= 8:   # the reference only ever runs align_avx for queryLength >= 8 (...SmithWaterman.cpp:62)
        av = o.ref_align(r, a, p, s, o.REF_AVX)
        assert sc == av, (r, a, p, s, sc, av)
    return sc


def mutate(rng, seq, alphabet, sub=0.03, ins=0.01, dele=0.01):
    out = []
    for ch in seq:
        x = rng.random()
        if x < dele:
            continue
        if x < dele + ins:
            out.append(rng.choice(alphabet))
        out.append(rng.choice(alphabet) if rng.random() < sub else ch)
    return "".join(out) or rng.choice(alphabet)


def fuzz_cases(n, seed):
    rng = random.Random(seed)
    alphabets = ["A", "AC", "ACG", "ACGT", "ACGTN", "acgtACGT", "AT"]
    for k in range(n):
        alpha = rng.choice(alphabets)
        tl = rng.choice([1, 2, 3, 5, 8, 15, 16, 17, 31, 32, 33, 47, 64, 65, 90, 129]) if k % 3 else rng.randint(1, 200)
        ref = "".join(rng.choice(alpha) for _ in range(tl))
        mode = rng.randrange(4)
        if mode == 0:                                   # independent random alt
            alt = "".join(rng.choice(alpha) for _ in range(rng.randint(1, 200)))
        elif mode == 1:                                 # exact substring
            i = rng.randrange(tl); j = rng.randint(i + 1, tl)
            alt = ref[i:j]
        elif mode == 2:                                 # mutated window + flanks
            i = rng.randrange(tl); j = rng.randint(i + 1, tl)
            alt = ("".join(rng.choice(alpha) for _ in range(rng.randint(0, 6))) + mutate(rng, ref[i:j], alpha)
                   + "".join(rng.choice(alpha) for _ in range(rng.randint(0, 6))))
        else:                                           # mutated full copy with a long indel
            alt = mutate(rng, ref, alpha, 0.02, 0.0, 0.0)
            if len(alt) > 12:
                c = rng.randrange(len(alt) - 8)
                alt = alt[:c] + alt[c + rng.randint(1, 8):] if rng.random() < 0.5 else alt[:c] + ref[:rng.randint(1, 8)] + alt[c:]
        yield ref, alt, SCORE_SETS[k % len(SCORE_SETS)]


def main():
    assert o.have_ref(), "build oracle/_ref first: make -C oracle ref"
    cases = []

    def add(kind, r, a, p):
        exp = [list(ref_all(r.encode("latin1"), a.encode("latin1"), p, s)) for s in o.STRATEGIES]
        cases.append({"kind": kind, "ref": r, "alt": a, "params": list(p), "expect": exp})

    for r, a in KATS:
        for p in (P_READ, P_HAP):
            add("kat", r, a, p)
    data = os.path.join(os.environ.get("MGL_REFERENCE", "/root/reference"), "src/test/resources/pairhmm-testdata.txt")
    for line in open(data):
        if line.startswith("#") or not line.strip():
            continue
        hap, read = line.split()[:2]
        for p in (P_READ, P_HAP):
            add("pairhmm", hap, read, p)
    for r, a, p in fuzz_cases(2500, 0x5157):
        add("fuzz", r, a, p)
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "sw_golden.json.gz")
    with gzip.GzipFile(out, "wb", mtime=0) as f:
        f.write(json.dumps({"strategies": list(o.STRATEGIES), "source": "reference align_scalar == align_avx == JNI entry",
                            "cases": cases}, separators=(",", ":")).encode())
    print(f"{len(cases)} cases x {len(o.STRATEGIES)} strategies -> {out} ({os.path.getsize(out)} bytes)")


if __name__ == "__main__":
    main()

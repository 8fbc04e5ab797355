#!/usr/bin/env python
"""Generate tests/golden/pairhmm_golden.json.gz from the REFERENCE ITSELF.

Runs in the authoring container only (needs /root/reference compiled into oracle/_ref/libmgl_pairhmm_ref.so by
`make -C oracle ref`).  The GPU box has no /root/reference; tests read only the committed fixture.

  file    the 104 cases of the reference's own fixture src/test/resources/pairhmm-testdata.txt with the expected
          log10 likelihoods the file carries (asserted at 1e-5 by MicrosoftPairHmmUnitTest.java:101), parsed the way
          that test parses them (quals - 33, base quals floored at 6), plus the output of the reference built here in
          float and in double mode, and the simpleTest known answer (MicrosoftPairHmmUnitTest.java:37-53)
  batch   seeded batches, every read against every haplotype through the reference's JNI entry
          (...MicrosoftPairHmm.cc:77): reads cut from the haplotypes with substitutions / indels (likelihoods near 1),
          unrelated reads (float underflow -> the double pass, .cc:176-203), 'N' bases on both sides

Domain.  The reference's AVX2 kernels, which its JNI entry uses on any AVX2 host, take the gap-continuation penalty
as ONE value per read (compute_prob_avxd.cc:482, fGapM) -- GATK passes a constant -- and hand batches containing a
haplotype shorter than the vector width to the scalar FLOAT kernel even in double mode (compute_prob_avxd.cc:524-528),
which returns values off by log10(2^900).  The batches stay inside the domain where the reference is self-consistent:
constant GCP per read, haplotypes of >= 8 bases.  Non-finite reference outputs are dropped.
"""
import gzip
import json
import math
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pairhmm_oracle as P  # noqa: E402

TESTDATA = "/root/reference/src/test/resources/pairhmm-testdata.txt"


def enc(b: bytes) -> str:
    return b.decode("latin1")


def make_batch(rng, n_reads, n_haps, alpha, related_frac):
    base = bytes(rng.choice(b"ACGT") for _ in range(rng.randint(120, 420)))
    haps = []
    for _ in range(n_haps):
        h = bytearray(base)
        for _ in range(rng.randint(0, 6)):                      # haplotypes of one region differ by a few edits
            k = rng.randrange(len(h))
            op = rng.random()
            if op < 0.6:
                h[k] = rng.choice(alpha)
            elif op < 0.8:
                del h[k]
            else:
                h.insert(k, rng.choice(alpha))
        haps.append(bytes(h))
    if rng.random() < 0.5:
        haps.append(bytes(rng.choice(alpha) for _ in range(rng.randint(8, 40))))     # a short one
    reads = []
    for _ in range(n_reads):
        n = rng.randint(1, 260)
        if rng.random() < related_frac:
            h = rng.choice(haps)
            n = min(n, len(h))
            s = rng.randrange(len(h) - n + 1)
            b = bytearray(h[s:s + n])
            for _ in range(rng.randint(0, 3)):
                b[rng.randrange(n)] = rng.choice(alpha)
            b = bytes(b)
        else:
            b = bytes(rng.choice(alpha) for _ in range(n))
        gcp = rng.choice((10, 10, 10, 6, 15, 20))
        reads.append((b, bytes(rng.randint(6, 45) for _ in range(n)), bytes(rng.randint(20, 50) for _ in range(n)),
                      bytes(rng.randint(20, 50) for _ in range(n)), bytes([gcp] * n)))
    return reads, haps


def main():
    assert P.have_ref(), "build oracle/_ref first: make -C oracle ref"
    out = {"file": [], "batch": []}
    for hap, read, expected in P.parse_testdata(open(TESTDATA).read()):
        out["file"].append({"hap": enc(hap), "read": [enc(x) for x in read], "expected_file": expected,
                            "ref_float": float(P.ref_likelihoods([read], [hap], False)[0, 0]),
                            "ref_double": float(P.ref_likelihoods([read], [hap], True)[0, 0])})
    q = b"++++"                                                   # raw bytes, not normalised, as simpleTest passes them
    out["simple"] = {"hap": "ACGT", "read": [enc(x) for x in (b"ACGT", q, q, q, q)], "expected": -6.022797e-01,
                     "ref_float": float(P.ref_likelihoods([(b"ACGT", q, q, q, q)], [b"ACGT"], False)[0, 0])}
    rng = random.Random(20261104)
    for k in range(8):
        alpha = b"ACGT" if k % 2 == 0 else b"ACGTN"
        reads, haps = make_batch(rng, 24, 6, alpha, related_frac=(0.9, 0.5, 0.1, 0.7)[k % 4])
        rf = P.ref_likelihoods(reads, haps, False)
        rd = P.ref_likelihoods(reads, haps, True)
        keep = [[bool(math.isfinite(rf[r, h]) and math.isfinite(rd[r, h])) for h in range(len(haps))] for r in range(len(reads))]
        out["batch"].append({"reads": [[enc(x) for x in r] for r in reads], "haps": [enc(h) for h in haps],
                             "ref_float": [[float(v) if math.isfinite(v) else None for v in row] for row in rf],
                             "ref_double": [[float(v) if math.isfinite(v) else None for v in row] for row in rd], "finite": keep})
    path = os.path.join(ROOT, "tests", "golden", "pairhmm_golden.json.gz")
    with gzip.open(path, "wb", 9) as f:
        f.write(json.dumps(out).encode())
    n = sum(len(b["reads"]) * len(b["haps"]) for b in out["batch"])
    print(f"{path}: {len(out['file'])} file cases, {n} batch pairs, {os.path.getsize(path)} bytes")


if __name__ == "__main__":
    main()

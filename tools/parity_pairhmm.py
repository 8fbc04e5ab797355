#!/usr/bin/env python
"""Full-job PairHMM parity: every (read, haplotype) pair of a synthetic region through the CUDA path (C-ABI, host
buffers) against the reference's own C++ (oracle/_ref: its JNI entry, AVX2 float + double escalation, or double
throughout) on all host cores.  Tolerance 1e-5 on the log10 likelihood (MicrosoftPairHmmUnitTest.java:101).
Test infrastructure: needs a B200 and the prebuilt oracle/_ref.  Prints one JSON line per mode (kept under profiles/).

  python tools/parity_pairhmm.py [--reads 8192 --haps 32 --unrelated 0.1]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=8192)
    ap.add_argument("--haps", type=int, default=32)
    ap.add_argument("--hap-len", type=int, default=400)
    ap.add_argument("--unrelated", type=float, default=0.1, help="fraction of reads replaced by random bases (float underflow -> double pass)")
    args = ap.parse_args()
    import bench_pairhmm as B
    from mgl_b200 import pairhmm
    from oracle import pairhmm_oracle as P
    if not P.have_ref():
        print(json.dumps({"error": "oracle/_ref not built"}))
        return 1
    lengths, rb, hb, cells = B.synth(args.reads, args.haps, 150, args.hap_len, seed=0xA11, mixed=True)
    rng = np.random.default_rng(7)
    acgtn = np.frombuffer(b"ACGTN", dtype=np.uint8)
    off = 0
    for r in range(args.reads):                       # unrelated reads, and a sprinkle of N in the others
        n = int(lengths[1 + r])
        if rng.random() < args.unrelated:
            rb[off:off + n] = acgtn[rng.integers(0, 4, n)]
        else:
            k = rng.integers(0, n, 1)
            if rng.random() < 0.2:
                rb[off + k] = ord("N")
        off += 5 * n
    hmm = pairhmm.MicrosoftPairHmm(0)
    hmm._ensure()
    rc = 0
    for use_double in (False, True):
        got = hmm.compute_buffers(lengths, rb, hb, None, use_double).reshape(args.reads, args.haps)
        st = dict(hmm.last_stats)
        secs, ref = P.ref_likelihoods_timed(lengths, rb, hb, use_double, 0, 8)
        ok = np.isfinite(ref)
        diff = np.abs(got - ref)
        worst = float(diff[ok].max())
        bad = int((diff[ok] > 1e-5).sum())
        print(json.dumps({"mode": "double" if use_double else "float + double escalation", "pairs": int(got.size), "cells": int(cells),
                          "checked": int(ok.sum()), "max_abs_diff_log10": worst, "over_tolerance": bad, "tolerance": 1e-5,
                          "pairs_recomputed_in_double_on_gpu": int(st["pairs_double"]), "gpu_launches": int(st["launches"]),
                          "non_finite_on_both_sides": bool(np.all(~np.isfinite(got[~ok]))),
                          "reference": f"oracle/_ref JNI entry, {P.ref_max_threads()} threads, {secs:.2f} s"}))
        rc |= 2 if bad else 0
    hmm.done()
    return rc


if __name__ == "__main__":
    sys.exit(main())

#!/usr/bin/env python
"""Full-size parity: every pair of a BASELINE.json config through the CUDA path (C-ABI,
host buffers) against the reference's own C++ (oracle/_ref: align_scalar or the JNI
dispatch rule) on all host cores.  Bit-exact CIGAR text and offset, or it reports the
first differing pairs.  Test infrastructure: needs a B200 and the prebuilt oracle/_ref.

  python tools/parity_full.py --config 2            # 1 M pairs
  python tools/parity_full.py --config 4 --pairs 200
  python tools/parity_full.py --config 5 --pairs 2000000 --every 1000   # first 1e6, then every 1000th

Prints one JSON line per config (kept under profiles/).
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

DEFAULT = {2: 1_000_000, 3: 100_000, 4: 10_000, 5: 1_000_000}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, required=True)
    ap.add_argument("--pairs", type=int, default=0)
    ap.add_argument("--every", type=int, default=1, help="beyond the first --dense pairs, check every k-th pair")
    ap.add_argument("--dense", type=int, default=1_000_000)
    ap.add_argument("--kind", default="dispatch", choices=["dispatch", "scalar", "avx"])
    ap.add_argument("--slice", type=int, default=0, help="pairs per reference call (0: sized for ~1 GB of CIGAR text)")
    ap.add_argument("--pack", type=int, default=-1, help="pack_arena option (2-bit transport of the arena): 0 / 1 / 2, -1 library default")
    args = ap.parse_args()

    import mgl_b200
    from mgl_b200 import synth
    from oracle import sw_oracle as O

    if not O.have_ref():
        print(json.dumps({"config": args.config, "error": "oracle/_ref not built"}))
        return 1
    n = args.pairs or DEFAULT[args.config]
    d = synth.generate(args.config, n)
    sw = mgl_b200.MicrosoftSmithWaterman(0)
    sw._ensure()
    if args.pack >= 0:
        sw.set_option("pack_arena", args.pack)
    total_len = int(d["ref_len"].astype(np.int64).sum() + d["alt_len"].astype(np.int64).sum())
    t0 = time.perf_counter()
    res = sw.align_arrays(d["arena"], d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"],
                          ops_cap=max(8 * n, total_len // 8))
    gpu_s = time.perf_counter() - t0

    idx = np.arange(n)
    if args.every > 1 and n > args.dense:
        idx = np.concatenate([idx[: args.dense], idx[args.dense:: args.every]])
    kind = {"dispatch": O.REF_DISPATCH, "scalar": O.REF_SCALAR, "avx": O.REF_AVX}[args.kind]
    threads = O.ref_max_threads()
    stride = int(4 * (int(d["ref_len"].max()) + int(d["alt_len"].max())) + 16)
    per = args.slice or max(64, min(len(idx), (1 << 30) // stride))
    bad, checked, ref_s = [], 0, 0.0
    cs = res.cigar_start.astype(np.int64)
    for lo in range(0, len(idx), per):
        sel = idx[lo: lo + per]
        secs, offs, texts = O.ref_align_batch(d["arena"], d["ref_off"][sel], d["alt_off"][sel], d["ref_len"][sel], d["alt_len"][sel],
                                              d["params"], strategies=d["strategies"][sel].astype("uint8"), threads=threads,
                                              kind=kind, want_output=True, cigar_stride=stride)
        ref_s += secs
        for k, p in enumerate(sel):
            o = res.ops[cs[p]: cs[p + 1]]
            got = "".join(f"{int(x) >> 4}{'MIDNSHP=X'[int(x) & 15]}" for x in o)
            if got != texts[k] or int(res.offsets[p]) != int(offs[k]):
                if len(bad) < 5:
                    bad.append({"pair": int(p), "tl": int(d["ref_len"][p]), "ql": int(d["alt_len"][p]),
                                "got": [got[:200], int(res.offsets[p])], "ref": [texts[k][:200], int(offs[k])]})
                elif len(bad) == 5:
                    bad.append("...")
            checked += 1
    mism = sum(1 for b in bad if b != "...")
    print(json.dumps({"config": args.config, "pairs": n, "checked": checked, "mismatches_shown": mism, "bit_exact": not bad,
                      "cells": int(d["cells"]), "reference": f"oracle/_ref {args.kind}, {threads} threads, {ref_s:.1f} s",
                      "gpu_e2e_s": round(gpu_s, 3), "pack_arena": args.pack, "h2d_bytes": int(res.stats["h2d_bytes"]), "pairs_s16": int(res.stats["pairs_s16"]), "pairs_s32": int(res.stats["pairs_s32"]),
                      "first_bad": bad}))
    sw.close()
    return 0 if not bad else 2


if __name__ == "__main__":
    sys.exit(main())

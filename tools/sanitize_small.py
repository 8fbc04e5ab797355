#!/usr/bin/env python
"""A small batch of every kernel shape through mgl_sw_align_batch, for compute-sanitizer:
  compute-sanitizer --tool memcheck python tools/sanitize_small.py
(test infrastructure; checks a sample against the oracle as well)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mgl_b200  # noqa: E402
from mgl_b200 import synth  # noqa: E402
from oracle import sw_oracle as O  # noqa: E402

sw = mgl_b200.MicrosoftSmithWaterman(0)
sw._ensure()
sw.set_option("chunk_pairs", 2048)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 6000
d = synth.generate(5, n)
keep = np.nonzero(d["alt_len"] < 3000)[0][:n]            # long pairs are slow under the sanitizer: keep a few
big = np.nonzero(d["alt_len"] >= 3000)[0][:2]
sel = np.sort(np.concatenate([keep, big]))
args = (d["arena"], d["ref_off"][sel], d["ref_len"][sel], d["alt_off"][sel], d["alt_len"][sel], d["params"], 1, d["strategies"][sel])
for hp in (0, 1):
    sw.set_option("host_plan", hp)
    res = sw.align_arrays(*args, ops_cap=64 * len(sel))
    sc, _ = sw.score_arrays(*args)
    bad = 0
    for k in range(0, len(sel), 37):
        p = int(sel[k])
        r, a = synth.pair_bytes(d, p)
        if len(r) * len(a) > 2_000_000:
            continue
        if (res.cigar(k), int(res.offsets[k])) != O.align(r, a, d["params"], int(d["strategies"][p])):
            bad += 1
    print("host_plan", hp, "pairs", len(sel), "chunks", res.stats["chunks"], "launches", res.stats["fill_launches"], "bad", bad)
    assert bad == 0
sw.close()
print("sanitize_small ok")

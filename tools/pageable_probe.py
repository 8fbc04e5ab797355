#!/usr/bin/env python
"""End-to-end time of one config-2 call from PAGEABLE caller memory (what a JVM hands over) for several counts of the
threads that stage the arena into pinned memory:  python tools/pageable_probe.py [threads ...]   (needs a B200)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import mgl_b200
from mgl_b200 import synth

n = 1_000_000
d = synth.generate(2, n, pinned=False)
arena = np.array(d["arena"], copy=True)
sw = mgl_b200.MicrosoftSmithWaterman(0); sw._ensure()
out = (np.zeros(n, np.int32), np.zeros(n + 1, np.uint64), np.zeros(8 * n, np.uint32))
args = (arena, d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"])
for T in [int(x) for x in sys.argv[1:]] or [8]:
    sw.set_option("copy_threads", T)
    for _ in range(3):
        sw.align_arrays(*args, out=out)
    t = time.perf_counter()
    for _ in range(5):
        sw.align_arrays(*args, out=out)
    ms = (time.perf_counter() - t) / 5 * 1e3
    print(f"copy_threads {T}: {ms:.2f} ms per 1 M pairs, {d['cells'] / ms / 1e6:.0f} GCUPS")

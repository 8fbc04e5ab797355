#!/usr/bin/env python
"""Latency and coalesced throughput of the reference-shaped one-pair entry (mgl_sw_align_one / JNI alignNative).
  python tools/one_pair_latency.py        (needs a B200)"""
import os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mgl_b200
from mgl_b200 import synth

d = synth.generate(2, 4096)
pairs = [synth.pair_bytes(d, p) for p in range(4096)]
P = (10, -15, -30, -5)
for r, a in pairs[:50]:
    mgl_b200.align_one_text(r, a, P, 1)
t = time.perf_counter()
for r, a in pairs[:1000]:
    mgl_b200.align_one_text(r, a, P, 1)
one = (time.perf_counter() - t) / 1000
print(f"1 thread: {one * 1e6:.1f} us per call, {1 / one:.0f} alignments/s")
for T in (4, 16, 64):
    def work(t):
        for k in range(t, len(pairs), T):
            mgl_b200.align_one_text(pairs[k][0], pairs[k][1], P, 1)
    th = [threading.Thread(target=work, args=(t,)) for t in range(T)]
    t0 = time.perf_counter()
    for x in th: x.start()
    for x in th: x.join()
    dt = time.perf_counter() - t0
    print(f"{T} threads: {len(pairs) / dt:.0f} alignments/s ({dt / len(pairs) * 1e6:.1f} us per call amortised)")

#!/usr/bin/env python
"""PairHMM over the GPUs of one box: reads are sharded (mgl_b200/shard.py pairhmm_shard_reads -- every (read, haplotype)
pair is independent, nothing is exchanged), one handle and one host thread per GPU in ONE process, host buffers in and
out through mgl_pairhmm_compute.  Weak scaling: 4 096 reads x 150 b per GPU against 32 haplotypes x 400 b.
One JSON line with the aggregate GCUPS at N = 1, 2, 4, ... visible GPUs."""
import importlib.util
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
spec = importlib.util.spec_from_file_location("bench_pairhmm", os.path.join(ROOT, "tools", "bench_pairhmm.py"))
bp = importlib.util.module_from_spec(spec)
spec.loader.exec_module(bp)
from mgl_b200 import pairhmm, shard  # noqa: E402

from mgl_b200 import native  # noqa: E402
ndev = native.lib().mgl_sw_device_count()
per_gpu, haps, read_len, hap_len, steps = 4096, 32, 150, 400, 5
rows = []
for n in [k for k in (1, 2, 4, 8) if k <= ndev]:
    lengths, rb, hb, cells = bp.synth(per_gpu * n, haps, read_len, hap_len)
    out = np.zeros(per_gpu * n * haps, dtype=np.float64)
    jobs = []
    for r in range(n):
        first, sub_lengths, sub_reads = shard.pairhmm_shard_reads(lengths, rb, n, r)
        h = pairhmm.MicrosoftPairHmm(r)
        h._ensure()
        jobs.append((h, np.ascontiguousarray(sub_lengths), np.ascontiguousarray(sub_reads), out[first * haps:(first + int(sub_lengths[0])) * haps]))

    def run(j):
        h, sl, sr, o = j
        h.compute_buffers(sl, sr, hb, o, False)

    def step():
        th = [threading.Thread(target=run, args=(j,)) for j in jobs[1:]]
        for t in th:
            t.start()
        run(jobs[0])
        for t in th:
            t.join()

    for _ in range(2):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    # the sharded job must equal the unsharded one on a sample
    ref = jobs[0][0].compute_buffers(*bp.sub_job(lengths, rb, 16), hb, None, False)
    ok = bool(np.allclose(ref, out[: 16 * haps], atol=1e-9))
    rows.append({"gpus": n, "reads": per_gpu * n, "pairs": per_gpu * n * haps, "cells": cells, "ms_per_step": dt * 1e3, "value": cells / dt / 1e9, "unit": "GCUPS",
                 "sharded_equals_unsharded_sample": ok})
    for h, *_ in jobs:
        h.done()
base = rows[0]["value"]
for r in rows:
    r["fraction_of_linear"] = r["value"] / (base * r["gpus"])
print(json.dumps({"what": "PairHMM, host buffers in and out, reads sharded over N GPUs by one process (weak scaling)", "rows": rows}))

#!/usr/bin/env python
"""Bench of the second kernel, PairHMM, in the shape of bench.py's JSON line.

  python tools/bench_pairhmm.py [--reads 4096 --haps 32 --read-len 150 --hap-len 400 --steps 5 --warmup 3]
  python tools/bench_pairhmm.py --impl reference        # the reference's own CPU PairHMM (oracle/_ref) on host cores

One synthetic "region": haplotypes that differ by a few edits, reads cut from them with sequencing-like errors and
qualities, every read against every haplotype (the job of computeLikelihoodsNative).  `value` = cells of the job
divided by the device time of its kernels with inputs resident in HBM (CUDA events on the engine's stream); `e2e` =
the same through mgl_pairhmm_compute with host buffers; `cpu_baseline` = the reference's AVX2 float path with its
double escalation, reads spread over all host cores, on a bounded sample of the same reads.
1 CUP = one cell (read row, haplotype column); algorithmic work per cell = 8 FP32 instructions = 11 flops
(compute_prob_scalar.cc:39-43: 3 fused multiply-adds, 3 multiplies, 2 adds).
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
METRIC = "GCUPS, PairHMM forward algorithm, float with double escalation (1 CUP = one read-row x haplotype-column cell)"
FLOPS_PER_CELL = 11.0


def synth(n_reads, n_haps, read_len, hap_len, seed=0xF3, mixed=False):
    rng = np.random.default_rng(seed)
    acgt = np.frombuffer(b"ACGT", dtype=np.uint8)
    base = acgt[rng.integers(0, 4, hap_len)]
    haps = []
    for _ in range(n_haps):
        h = base.copy()
        for k in rng.integers(0, hap_len, 4):
            h[k] = acgt[rng.integers(0, 4)]
        haps.append(h)
    reads = []
    lens = []
    for _ in range(n_reads):
        if mixed:                                     # a region of trimmed and untrimmed reads: one launch per shape
            read_len = int(rng.choice((76, 101, 101, 150, 150, 150, 250)))
        lens.append(read_len)
        h = haps[int(rng.integers(0, n_haps))]
        s = int(rng.integers(0, hap_len - read_len + 1))
        b = h[s:s + read_len].copy()
        err = rng.random(read_len) < 0.01
        b[err] = acgt[rng.integers(0, 4, int(err.sum()))]
        q = np.clip(rng.normal(32, 6, read_len), 6, 41).astype(np.uint8)
        q[err] = 8
        reads.append(np.concatenate([b, q, np.full(read_len, 45, np.uint8), np.full(read_len, 45, np.uint8), np.full(read_len, 10, np.uint8)]))
    lengths = np.array([n_reads] + lens + [n_haps] + [hap_len] * n_haps, dtype=np.int32)
    return lengths, np.concatenate(reads), np.concatenate(haps), int(sum(lens)) * n_haps * hap_len


def sub_job(lengths, rb, n_reads_sample):
    R = int(lengths[0])
    n = min(R, n_reads_sample)
    rl = lengths[1:1 + n]
    sub = np.concatenate([[n], rl, lengths[1 + R:]]).astype(np.int32)
    return sub, rb[: 5 * int(rl.sum())]


def cpu_reference(lengths, rb, hb, budget_s=12.0):
    from oracle import pairhmm_oracle as P
    if not P.have_ref():
        return None
    threads = P.ref_max_threads()
    R = int(lengths[0]); H = int(lengths[1 + R])
    hl = int(lengths[2 + R])
    n = 64
    while True:
        sub, srb = sub_job(lengths, rb, n)
        secs, _ = P.ref_likelihoods_timed(sub, srb, hb, False, threads, 8)
        if secs >= budget_s / 4 or n >= R:
            break
        n = min(R, int(n * max(2.0, budget_s / 2 / max(secs, 1e-3))))
    cells = int(sub[1:1 + int(sub[0])].sum()) * H * hl
    return {"value": cells / secs / 1e9, "unit": "GCUPS", "cores": threads, "kind": "reference",
            "sample": f"first {int(sub[0])} of {R} reads x {H} haplotypes, oracle/_ref (the reference's AVX2 float kernel + double escalation behind its JNI entry), "
                      f"8 reads per JNI call over {threads} OpenMP threads, {secs:.2f} s"}, secs


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=4096)
    ap.add_argument("--haps", type=int, default=32)
    ap.add_argument("--read-len", type=int, default=150)
    ap.add_argument("--hap-len", type=int, default=400)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--double", action="store_true", help="double throughout (useDoublePrecision)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--mixed", action="store_true", help="read lengths drawn from 76 / 101 / 150 / 250")
    args = ap.parse_args()
    lengths, rb, hb, cells = synth(args.reads, args.haps, args.read_len, args.hap_len, mixed=args.mixed)
    workload = f"{args.reads} reads x {'76-250 (mixed)' if args.mixed else args.read_len} b against {args.haps} haplotypes x {args.hap_len} b (one region, every pair)"

    if args.impl == "reference":
        cb, secs = cpu_reference(lengths, rb, hb, budget_s=20.0)
        line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": "GCUPS", "n_gpus": 1, "steps": 1, "warmup": 0,
                "ms_per_step": secs * 1e3, "higher_is_better": True, "dtype": "f32 / f64", "data": "synthetic",
                "config": {"workload": workload}, "cpu_baseline": cb,
                "e2e": {"value": cb["value"], "unit": "GCUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return 0

    import bench
    from mgl_b200 import pairhmm
    hmm = pairhmm.MicrosoftPairHmm(0)
    hmm._ensure()
    sampler = bench.ClockSampler(0)
    sampler.start()
    # device-resident: inputs uploaded once, kernels timed with CUDA events on the engine's stream
    hmm.resident_prepare(lengths, rb, hb)
    for _ in range(args.warmup):
        hmm.resident_run(args.double)
    dev_ms, launches, redone = 0.0, 0, 0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        st = hmm.resident_run(args.double)
        dev_ms += st["device_ms"]; launches += st["launches"]; redone = st["pairs_double"]
    wall_ms = (time.perf_counter() - t0) * 1e3
    # e2e: host buffers in, host doubles out, through the C-ABI
    out = np.zeros(args.reads * args.haps, dtype=np.float64)
    for _ in range(min(args.warmup, 2)):
        hmm.compute_buffers(lengths, rb, hb, out, args.double)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        hmm.compute_buffers(lengths, rb, hb, out, args.double)
    e2e_ms = (time.perf_counter() - t0) * 1e3
    st = hmm.last_stats
    clocks = sampler.stop()
    K = args.steps
    value = cells * K / (dev_ms * 1e-3) / 1e9
    sm_mhz = clocks.get("sm_mhz") or 1965.0
    peak = 148 * 128 * 2 * sm_mhz * 1e6 / 1e12
    line = {"metric": METRIC, "value": value, "unit": "GCUPS", "n_gpus": 1, "steps": K, "warmup": args.warmup,
            "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64" if args.double else "f32 (f64 for the escalated pairs)", "data": "synthetic",
            "config": {"workload": workload, "pairs": args.reads * args.haps, "cells": cells, "pairs_recomputed_in_double": int(redone),
                       "l2": "no O(cells) memory traffic: a pair reads 5 x readLen + hapLen bytes and writes 8"},
            "wall_ms_per_step": wall_ms / K, "gpu_launches": int(launches), "clocks": clocks,
            "roofline": {"bound": "fp32", "achieved": value * FLOPS_PER_CELL / 1e3, "peak": peak, "unit": "TFLOP/s",
                         "frac": value * FLOPS_PER_CELL / 1e3 / peak, "traffic": None,
                         "note": "dominant kernel = ph_forward_kernel<float>; algorithmic 11 flops per cell (8 FP32 instructions) x cells / device time; "
                                 "peak = 148 SMs x 128 FP32 lanes x 2 x the SM clock sampled during the run (nominal; MEASURED_PEAKS.json has no FP32 figure)"},
            "e2e": {"value": cells * K / (e2e_ms * 1e-3) / 1e9, "unit": "GCUPS", "h2d_bytes_per_step": int(st["h2d_bytes"]),
                    "d2h_bytes_per_step": int(st["d2h_bytes"]), "ms_per_step": e2e_ms / K}}
    if not args.no_cpu_baseline:
        cb = cpu_reference(lengths, rb, hb)
        if cb:
            line["cpu_baseline"] = cb[0]
    # parity of this very job against the reference on a sample (not timed)
    try:
        from oracle import pairhmm_oracle as P
        if P.have_ref():
            sub, srb = sub_job(lengths, rb, 64)
            got = hmm.compute_buffers(sub, srb, hb, None, args.double).reshape(int(sub[0]), args.haps)
            ref = P.ref_likelihoods_timed(sub, srb, hb, args.double)[1]
            line["parity_sample"] = {"pairs": int(got.size), "max_abs_diff_log10": float(np.max(np.abs(got - ref))), "tolerance": 1e-5}
    except Exception as e:  # noqa: BLE001
        line["parity_sample"] = {"error": str(e)}
    hmm.done()
    print(json.dumps(line))
    return 0


if __name__ == "__main__":
    sys.exit(main())

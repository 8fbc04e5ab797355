#!/usr/bin/env python
"""bench.py -- the headline benchmark: Smith-Waterman (fill + traceback + CIGAR)
throughput on BASELINE.json config 2 (1 M synthetic 150 bp reads vs 300-500 bp
haplotypes, scores (10,-15,-30,-5), SOFTCLIP) per GPU.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--pairs P] [--config C]

A step is one pass of the hot path over the whole batch.  `value` is GCUPS with
inputs resident in HBM (CUDA events on the launching streams around the whole
batch, max over ranks); `e2e` is the same metric through the C-ABI
mgl_sw_align_batch with pinned HOST buffers in and out (`e2e_pageable`: the same
call from ordinary pageable memory, what a JVM direct ByteBuffer is).  N > 1: one
process per GPU (torchrun), each rank aligns its own shard of the synthetic
stream, no data-path collective (pairs are independent) -> weak scaling.

The N = 1 line also carries `configs`: BASELINE.json configs 3, 4 and 5 measured
the same way (3 steps each), and `pairhmm`, the library's second kernel; the
N > 1 line carries `config5` (the scaling sweep's workload, 1 M mixed pairs per
GPU standing in for 100 M: at ~1.4 TCUPS the full 1.5e13 cells are ~3 GPU-hours)
and `multi_handle`: ONE process driving all N GPUs through mgl_sw_create_multi.

--impl reference times the reference's own CPU implementation (oracle/_ref:
the unmodified align_avx / align_scalar behind the JNI dispatch rule, every CPU
this process may run on) on the same pairs per step, rank 0 only.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "GCUPS, Smith-Waterman affine-gap fill + traceback + CIGAR (1 CUP = one DP cell)"
UNIT = "GCUPS"
ALGO_OPS_PER_CELL = 10          # SURVEY.md 8(d), counted from sw.cpp:55-93
SUB_STEPS, SUB_WARMUP = 3, 3    # the `configs` / `config5` sub-records


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", type=int, default=2, help="BASELINE.json config index (2, 3, 4 or 5)")
    ap.add_argument("--pairs", type=int, default=0, help="pairs per GPU (default: the config's size)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-sub", action="store_true", help="skip the configs / config5 / pairhmm / multi_handle sub-records")
    ap.add_argument("--no-numa-bind", action="store_true", help="N > 1: do not bind each rank to the CPUs local to its GPU")
    ap.add_argument("--chunk-pairs", type=int, default=0, help="engine chunk size (0: library default)")
    ap.add_argument("--pipe-wide", type=int, default=-1, help="long-pair kernel on eight warps per pair: 0 never, 1 auto, 2 always (-1: library default)")
    ap.add_argument("--pack", type=int, default=-1, help="2-bit transport of all-ACGT arena ranges: 0 never, 1 when the planner is ahead, 2 always (-1: library default)")
    ap.add_argument("--pack-lead", type=int, default=-1, help="chunks a planner must be ahead of the upload to pack (-1: library default)")
    ap.add_argument("--plan-threads", type=int, default=0, help="host planner threads (0: library default)")
    ap.add_argument("--slots", type=int, default=-1, help="work slots (chunks in flight): 0 auto, 2..4 (-1: library default)")
    ap.add_argument("--host-plan", type=int, default=-1, help="1: plan chunks on the host (the round-1 path) instead of on the device (-1: library default)")
    return ap.parse_args()


DEFAULT_PAIRS = {2: 1_000_000, 3: 100_000, 4: 1_000, 5: 1_000_000}
SUB_PAIRS = {3: 100_000, 4: 4_000, 5: 1_000_000}
WORKLOAD = {2: "config 2: 150 bp reads vs 300-500 bp haplotypes, (10,-15,-30,-5), SOFTCLIP",
            3: "config 3: 300-1000 bp hap-to-ref, (200,-150,-260,-11), SOFTCLIP/INDEL",
            4: "config 4: 5-10 kbp pairs, (200,-150,-260,-11), SOFTCLIP",
            5: "config 5: mixed 90 % c2 / 9.9 % c3 / 0.1 % c4 shapes"}


def config_dict(config, n):
    """The workload both arms name, key for key (the driver compares them)."""
    return {"workload": WORKLOAD[config], "pairs_per_gpu": n,
            "l2": "inputs_exceed_l2 (arena + checkpoints per step >> 126 MB)", "sharding": "independent pairs, no collective"}


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.gpu, self.rows, self.proc = gpu, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2])); pw.append(float(r[3]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def dist_setup(n_gpus):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        global _CPU_GROUP
        _CPU_GROUP = dist.new_group(backend="gloo")      # a barrier that keeps the waiting ranks off the GPUs (multi_handle)
        return rank, world, local, dist
    return rank, world, local, None


_CPU_GROUP = None


def cpu_barrier(dist):
    """ranks wait in a socket, not in an NCCL kernel: nothing of theirs runs on any GPU meanwhile"""
    if dist is not None:
        dist.barrier(group=_CPU_GROUP)


def reduce_max_sum(dist, local, tmax, tsum):
    """max over ranks of each value in tmax, sum over ranks of each in tsum (device tensors over NCCL)."""
    if dist is None:
        return tmax, tsum
    import torch
    a = torch.tensor(tmax, dtype=torch.float64, device=f"cuda:{local}")
    b = torch.tensor(tsum, dtype=torch.float64, device=f"cuda:{local}")
    dist.all_reduce(a, op=dist.ReduceOp.MAX)
    dist.all_reduce(b, op=dist.ReduceOp.SUM)
    return a.tolist(), b.tolist()


def barrier(dist, local):
    if dist is not None:
        import torch
        dist.barrier(device_ids=[local])
        torch.cuda.synchronize(local)


# ---------------------------------------------------------------- the reference's CPU path (oracle/_ref)

def cpu_reference_run(d, n_sample, threads, want_output=False):
    """Time the reference's own CPU path (oracle/_ref) on the first n_sample pairs."""
    from oracle import sw_oracle as O
    sl = slice(0, n_sample)
    secs, _, _ = O.ref_align_batch(d["arena"], d["ref_off"][sl], d["alt_off"][sl], d["ref_len"][sl], d["alt_len"][sl],
                                   d["params"], strategies=d["strategies"][sl].astype("uint8"), threads=threads,
                                   kind=O.REF_DISPATCH, want_output=want_output)
    cells = int((d["ref_len"][sl].astype("uint64") * d["alt_len"][sl].astype("uint64")).sum())
    return secs, cells


def cpu_baseline(d, budget_s=12.0):
    from oracle import sw_oracle as O
    if not O.have_ref():
        return None
    threads = O.ref_max_threads()                      # the CPUs this process may run on (not OMP_NUM_THREADS)
    n = len(d["ref_len"])
    probe = min(n, max(64, 40 * threads))
    secs, cells = cpu_reference_run(d, probe, threads)
    rate = probe / max(secs, 1e-6)
    n_sample = int(min(n, max(probe, rate * budget_s)))
    secs, cells = cpu_reference_run(d, n_sample, threads)
    # the 1-core figure SURVEY.md 8d asks for beside the all-core one: a sample sized for ~2 s on one thread
    n_one = int(min(n, max(16, n_sample / max(secs, 1e-6) / max(threads, 1) * 2.0)))
    secs1, cells1 = cpu_reference_run(d, n_one, 1)
    return {"value": cells / secs / 1e9, "unit": UNIT, "alignments_per_s": n_sample / secs, "cores": threads,
            "kind": "reference", "sample": f"first {n_sample} pairs of the workload, {cells / 1e9:.2f} Gcells, {secs:.2f} s, "
            f"reference align_avx/align_scalar via the JNI dispatch rule, one pair per OpenMP task",
            "one_core": {"value": cells1 / secs1 / 1e9, "unit": UNIT, "sample": f"first {n_one} pairs, {secs1:.2f} s on one thread"}}


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation on the same pairs per step, rank 0 only.  Maps
    oracle/ libraries only (the reference, and oracle/'s build of the input generator) -- nothing of the product."""
    if rank != 0:
        return
    from mgl_b200 import synth
    from oracle import sw_oracle as O
    O.build()
    synth.use_library(O.SYNTH_SO)           # same generator source, oracle/'s own build of it
    n = args.pairs or DEFAULT_PAIRS[args.config]
    if not O.have_ref():
        emit(({"impl": "reference", "unavailable": "oracle/_ref/libmgl_sw_ref.so was not built (no reference sources on this box)"}))
        return
    d = synth.generate(args.config, n)
    threads = O.ref_max_threads()
    # every step = the whole per-GPU workload, as in our arm; only where that cannot end within a few minutes
    # (config 4's long pairs: 56 alignments/s on 16 cores) a step is the first n_sample pairs of it
    probe = min(n, max(64, 40 * threads))
    secs, _ = cpu_reference_run(d, probe, threads)
    per_step_budget = 240.0 / max(1, args.steps + args.warmup)
    n_sample = int(min(n, max(probe, probe / max(secs, 1e-6) * per_step_budget)))
    if n_sample > 0.8 * n:
        n_sample = n
    for _ in range(args.warmup):
        cpu_reference_run(d, n_sample, threads)
    tot_s, tot_cells = 0.0, 0
    for _ in range(args.steps):
        s, c = cpu_reference_run(d, n_sample, threads)
        tot_s += s
        tot_cells += c
    val = tot_cells / tot_s / 1e9
    sample = (f"each step = all {n} pairs of the per-GPU workload" if n_sample == n else
              f"each step = first {n_sample} of the {n} pairs of the per-GPU workload") + f" ({tot_cells / args.steps / 1e9:.2f} Gcells)"
    emit(({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": tot_s / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        "config": config_dict(args.config, n),
        "alignments_per_s": n_sample * args.steps / tot_s,
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "reference", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


_REAL_STDOUT = None


def bind_to_gpu_cpus(gpu):
    """One process per GPU: run this rank (and allocate its pinned buffers, first touch) on the CPUs NVML reports as
    local to its GPU, so that host-to-device copies do not cross the socket interconnect.  Application-side policy (the
    library never changes its caller's affinity).  Returns the CPU count bound to, or None if NVML says nothing."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(gpu)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w in range(words) for b in range(64) if (int(mask[w]) >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus and len(cpus) < len(os.sched_getaffinity(0)):
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None


def protect_stdout():
    """Libraries (NCCL's version banner, torchrun) write to fd 1; the contract is ONE JSON line there.
    Everything else goes to stderr; emit() writes the line to the real stdout."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def emit(obj):
    _REAL_STDOUT.write(json.dumps(obj) + "\n")
    _REAL_STDOUT.flush()


# ---------------------------------------------------------------- our arm: one workload, measured three ways

def measure_device(sw, d, steps, warmup, dist, local):
    """Inputs resident in HBM, kernels timed with CUDA events on their streams.
    Pass 1 (`value`): the engine's production schedule -- chunks rotate through the work slots, so the traceback of
    chunk c overlaps the fill of the next chunks; one event pair brackets the whole batch.
    Pass 2 (roofline of the dominant kernel): one chunk at a time, every phase bracketed by events on its stream."""
    sw.resident_prepare(d["arena"], d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"])
    sw.set_option("overlap", 1)
    for _ in range(warmup):
        sw.resident_run()
    barrier(dist, local)
    r = {"dev_ms": 0.0, "launches": 0}
    t0 = time.perf_counter()
    for _ in range(steps):
        st = sw.resident_run()
        r["dev_ms"] += st["device_ms"]
        r["launches"] += st["fill_launches"] + st["other_launches"]
    barrier(dist, local)
    r["wall_ms"] = (time.perf_counter() - t0) * 1e3
    r["cells_step"] = st["cells"]
    r["all_s16"] = st["pairs_s32"] == 0 and st["pairs_s16"] > 0     # every pair in u16x2 arithmetic: two cells per lane-op
    r["pairs_s16"], r["pairs_s32"] = st["pairs_s16"], st["pairs_s32"]
    r["ckpt_bytes_step"] = st["trace_bytes"]           # checkpoint bytes the fill launches of one step write (planner's count)
    sw.set_option("overlap", 0)
    r.update(fill_ms=0.0, trace_ms=0.0, post_ms=0.0, serial_ms=0.0, fill_launches=0)
    for k in range(min(warmup, 2) + steps):
        st = sw.resident_run()
        if k >= min(warmup, 2):
            r["fill_ms"] += st["fill_ms"]; r["trace_ms"] += st["trace_ms"]; r["post_ms"] += st["post_ms"]
            r["serial_ms"] += st["device_ms"]; r["fill_launches"] += st["fill_launches"]
    sw.set_option("overlap", 1)
    sw.resident_drop()
    return r


def ops_capacity(d):
    import numpy as np
    # CIGAR element capacity: 8 per pair covers the short-read configs; long pairs carry ~1 element per 40 bases
    total_len = int(d["ref_len"].astype(np.int64).sum() + d["alt_len"].astype(np.int64).sum())
    return max(8 * len(d["ref_len"]), total_len // 8)


def measure_e2e(sw, d, steps, warmup, dist, local, pinned=True):
    """mgl_sw_align_batch, host buffers in and out, copies inside the timed region.  pinned=False: the caller's
    arena and result arrays are ordinary pageable memory (a JVM direct ByteBuffer)."""
    import numpy as np
    from mgl_b200 import native
    n = len(d["ref_len"])
    alloc = native.pinned_array if pinned else (lambda shape, dt: np.zeros(shape, dtype=dt))
    out = (alloc(n, np.int32), alloc(n + 1, np.uint64), alloc(ops_capacity(d), np.uint32))
    arena = d["arena"] if pinned else np.array(d["arena"], copy=True)
    call = lambda: sw.align_arrays(arena, d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"], out=out)
    for _ in range(warmup):
        call()
    barrier(dist, local)
    t0, c0 = time.perf_counter(), time.process_time()
    for _ in range(steps):
        r = call()
    ms, cpu_s = (time.perf_counter() - t0) * 1e3, time.process_time() - c0
    barrier(dist, local)
    return ms, int(r.stats["h2d_bytes"]), int(r.stats["d2h_bytes"]), cpu_s


def measure_score_only(sw, d, steps, dist, local):
    import numpy as np
    from mgl_b200 import native
    n = len(d["ref_len"])
    so_out = native.pinned_array(3 * n, np.int32).reshape(n, 3)
    call = lambda: sw.score_arrays(d["arena"], d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"], out=so_out)
    call()
    barrier(dist, local)
    t0 = time.perf_counter()
    for _ in range(steps):
        call()
    ms = (time.perf_counter() - t0) * 1e3
    barrier(dist, local)
    return ms


def rooflines(r, d, K, p_int32, hbm_peak, measured):
    """roofline / roofline_hbm of the fill launches of one workload from measure_device()'s per-kernel pass."""
    fill_cups = r["cells_step"] * K / (r["fill_ms"] * 1e-3)
    # the denominator in the arithmetic the kernel runs: a u16x2 lane-op is two 16-bit cell-ops
    p_int = (2.0 * p_int32 if r["all_s16"] else p_int32) if p_int32 else None
    launches_step = max(1, r["fill_launches"] // K)
    arena_bytes = int(d["arena"].nbytes)
    ach = fill_cups * ALGO_OPS_PER_CELL / 1e9
    roof = {"bound": "int", "achieved": ach, "peak": p_int, "unit": "Gop/s", "frac": (ach / p_int) if p_int else None,
            "traffic": (r["ckpt_bytes_step"] + arena_bytes) / launches_step,
            "note": "dominant kernel = sw_fill_kernel (all its launches of a step, timed with CUDA events on their stream in the "
                    "one-chunk-at-a-time pass); algorithmic 10 int ops/cell (SURVEY.md 8d) x cells; peak = max(iadd3, imnmx) measured "
                    "live by int_peak.cu (not in MEASURED_PEAKS.json)"
                    + (", x 2 because every pair of this workload runs in u16x2 arithmetic (one lane-op = two 16-bit cell-ops); "
                       "against the plain 32-bit peak (peak_s32) the fraction would be twice this" if r["all_s16"] else
                       "; pairs that run in u16x2 arithmetic do two cells per lane-op, so a mixed workload can exceed its 32-bit fraction")
                    + "; traffic = bytes counted in THIS run per average fill launch: the checkpoint bytes the planner laid out "
                      "(= what the launches write; ncu dram__bytes_write agrees within 1 %, profiles/) + the arena bytes they read once",
            "peak_s32": p_int32}
    hbm = {"bound": "hbm", "achieved": r["ckpt_bytes_step"] * K / (r["fill_ms"] * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
           "frac": r["ckpt_bytes_step"] * K / (r["fill_ms"] * 1e-3) / 1e9 / hbm_peak,
           "algorithmic_bytes_per_cell": 0.5, "checkpoint_bytes_per_cell": r["ckpt_bytes_step"] / r["cells_step"],
           "note": "checkpoint writes of the fill kernel (the only O(cells) HBM traffic on the path) against MEASURED_PEAKS.json hbm_gbs"
                   + ("" if measured else " (fallback 6650)") + "; algorithmic = SURVEY.md 8d's 4-bit trace"}
    return roof, hbm


def sub_config(sw, config, n, rank, dist, local, p_int32, hbm_peak, measured, want_e2e=True):
    """One BASELINE.json config as a sub-record: value, fill ms, roofline fraction, checkpoint B/cell, e2e."""
    from mgl_b200 import synth
    d = synth.generate(config, n, first=rank * n, pinned=True, threads=4)
    r = measure_device(sw, d, SUB_STEPS, SUB_WARMUP, dist, local)
    e2e_ms = measure_e2e(sw, d, SUB_STEPS, 1, dist, local)[0] if want_e2e else 0.0
    (dev_ms, fill_ms, e2e_ms), (cells_all, pairs_all) = reduce_max_sum(
        dist, local, [r["dev_ms"], r["fill_ms"], e2e_ms], [float(r["cells_step"]), float(n)])
    if rank != 0:
        return None
    K = SUB_STEPS
    roof, hbm = rooflines(r, d, K, p_int32, hbm_peak, measured)
    rec = {"workload": WORKLOAD[config], "pairs_per_gpu": n, "cells_per_gpu": r["cells_step"], "steps": K, "warmup": SUB_WARMUP,
           "value": cells_all * K / (dev_ms * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": dev_ms / K,
           "fill_ms_per_step": r["fill_ms"] / K, "traceback_ms_per_step": r["trace_ms"] / K, "fill_launches": r["fill_launches"] // K,
           "pairs_s16": r["pairs_s16"], "pairs_s32": r["pairs_s32"],
           "frac": roof["frac"], "roofline_peak_gops": roof["peak"], "checkpoint_bytes_per_cell": hbm["checkpoint_bytes_per_cell"],
           "hbm_frac": hbm["frac"], "traffic": roof["traffic"]}
    if want_e2e:
        rec["e2e"] = {"value": cells_all * K / (e2e_ms * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": e2e_ms / K}
    return rec


def pairhmm_record(local, ffma_gops, steps=5, warmup=3):
    """The library's second kernel on one synthetic region: 4 096 reads x 150 b against 32
    haplotypes x 400 b, every (read, haplotype) pair.  Device-resident value, e2e through mgl_pairhmm_compute, FP32
    roofline against the FMA rate measured on this box (int_peak.cu ffma_f32 x 2 flops)."""
    import importlib.util
    import numpy as np
    spec = importlib.util.spec_from_file_location("bench_pairhmm", os.path.join(ROOT, "tools", "bench_pairhmm.py"))
    bp = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bp)
    from mgl_b200 import pairhmm
    reads, haps, read_len, hap_len = 4096, 32, 150, 400
    lengths, rb, hb, cells = bp.synth(reads, haps, read_len, hap_len)
    hmm = pairhmm.MicrosoftPairHmm(local)
    hmm._ensure()
    hmm.resident_prepare(lengths, rb, hb)
    for _ in range(warmup):
        hmm.resident_run(False)
    dev_ms = 0.0
    for _ in range(steps):
        st = hmm.resident_run(False)
        dev_ms += st["device_ms"]
    redone = st["pairs_double"]
    out = np.zeros(reads * haps, dtype=np.float64)
    hmm.compute_buffers(lengths, rb, hb, out, False)
    t0 = time.perf_counter()
    for _ in range(steps):
        hmm.compute_buffers(lengths, rb, hb, out, False)
    e2e_ms = (time.perf_counter() - t0) * 1e3
    hmm.done()
    value = cells * steps / (dev_ms * 1e-3) / 1e9
    peak = 2.0 * ffma_gops / 1e3 if ffma_gops else None                  # TFLOP/s
    ach = value * bp.FLOPS_PER_CELL / 1e3
    return {"workload": f"{reads} reads x {read_len} b against {haps} haplotypes x {hap_len} b (one region, every pair)",
            "pairs": reads * haps, "cells": cells, "steps": steps, "value": value, "unit": UNIT, "ms_per_step": dev_ms / steps,
            "pairs_recomputed_in_double": int(redone),
            "roofline": {"bound": "fp32", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": (ach / peak) if peak else None,
                         "peak_nominal": 148 * 128 * 2 * 1.965e9 / 1e12,
                         "note": "ph_forward_kernel<float>; algorithmic 11 flops per cell (8 FP32 instructions, compute_prob_scalar.cc:39-43) "
                                 "x cells / device time; peak = 2 flops x the FP32 FMA issue rate measured on this box by int_peak.cu (ffma_f32)"},
            "e2e": {"value": cells * steps / (e2e_ms * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": e2e_ms / steps}}


def multi_handle_record(world, n_per_gpu, config=2):
    """ONE process, one handle over all `world` GPUs (mgl_sw_create_multi): the JVM's situation -- a single caller
    handing a batch of world x n pairs to the library, which cuts it by DP cells, runs one engine per GPU on its own
    host thread and gathers in input order.  Host buffers in and out, like `e2e`."""
    import numpy as np
    import mgl_b200
    from mgl_b200 import native, synth
    n = world * n_per_gpu
    d = synth.generate(config, n, pinned=True, threads=16)
    out = (native.pinned_array(n, np.int32), native.pinned_array(n + 1, np.uint64), native.pinned_array(ops_capacity(d), np.uint32))
    m = mgl_b200.MicrosoftSmithWaterman(list(range(world)))
    m._ensure()
    call = lambda: m.align_arrays(d["arena"], d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"], out=out)
    for _ in range(2):
        call()
    steps = 3
    t0, c0 = time.perf_counter(), time.process_time()
    for _ in range(steps):
        r = call()
    ms, cpu_s = (time.perf_counter() - t0) * 1e3, time.process_time() - c0
    m.close()
    return {"devices": world, "pairs": n, "cells": int(d["cells"]), "steps": steps, "value": d["cells"] * steps / (ms * 1e-3) / 1e9, "unit": UNIT,
            "ms_per_step": ms / steps, "h2d_bytes_per_step": int(r.stats["h2d_bytes"]), "d2h_bytes_per_step": int(r.stats["d2h_bytes"]),
            "host_cpu_ns_per_pair": cpu_s / (n * steps) * 1e9,
            "note": "mgl_sw_align_batch on a handle made by mgl_sw_create_multi over every GPU of the run, called from rank 0 alone while "
                    "the other ranks idle at a barrier; pinned host buffers in and out"}


def main():
    args = parse()
    protect_stdout()
    rank, world, local, dist = dist_setup(args.gpus)
    if args.impl == "reference":
        run_reference(args, rank, world)
        if dist is not None:
            dist.destroy_process_group()
        return

    import numpy as np
    import mgl_b200
    from mgl_b200 import native, synth

    numa = bind_to_gpu_cpus(local) if (world > 1 and not args.no_numa_bind) else None
    n = args.pairs or DEFAULT_PAIRS[args.config]
    d = synth.generate(args.config, n, first=rank * n, pinned=True, threads=4)       # this rank's shard of the stream
    sw = mgl_b200.MicrosoftSmithWaterman(local)
    sw._ensure()                                                          # no GPU / no extension -> raises
    for name, v in (("chunk_pairs", args.chunk_pairs), ("plan_threads", args.plan_threads)):
        if v:
            sw.set_option(name, v)
    for name, v in (("pipe_wide", args.pipe_wide), ("pack_arena", args.pack), ("pack_lead", args.pack_lead), ("host_plan", args.host_plan), ("slots", args.slots)):
        if v >= 0:
            sw.set_option(name, v)

    # integer issue rates (roofline denominator), measured on this box
    peaks = {}
    if rank == 0:
        L = native.lib()
        import ctypes as C
        for k in range(L.mgl_sw_int_peak_count()):
            ops, mhz = C.c_double(), C.c_double()
            if L.mgl_sw_int_peak(local, k, C.byref(ops), C.byref(mhz)) == 0:
                peaks[L.mgl_sw_int_peak_name(k).decode()] = ops.value / 1e9
    p_int32 = max(peaks.get("iadd3_s32", 0.0), peaks.get("imnmx_s32", 0.0)) or None
    measured = {}
    try:
        measured = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = measured.get("hbm_gbs", 6650.0)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    r = measure_device(sw, d, args.steps, args.warmup, dist, local)

    # ---- e2e: host buffers through the C-ABI, copies inside the timed region; then the score-only entry
    e2e_ms = so_ms = pg_ms = e2e_cpu_s = 0.0
    h2d = d2h = 0
    if not args.no_e2e:
        e2e_ms, h2d, d2h, e2e_cpu_s = measure_e2e(sw, d, args.steps, min(args.warmup, 2), dist, local)
        so_ms = measure_score_only(sw, d, args.steps, dist, local)
    clocks = sampler.stop() if rank == 0 else None     # sampled over the timed passes (value, per-kernel, e2e)
    if not args.no_e2e:
        pg_ms = measure_e2e(sw, d, min(args.steps, 5), 1, dist, local, pinned=False)[0]
    pg_steps = min(args.steps, 5)

    (dev_ms, fill_ms, wall_ms, e2e_ms, so_ms, pg_ms), (cells_all, pairs_all) = reduce_max_sum(
        dist, local, [r["dev_ms"], r["fill_ms"], r["wall_ms"], e2e_ms, so_ms, pg_ms], [float(r["cells_step"]), float(n)])

    line = None
    if rank == 0:
        K = args.steps
        cells_step = r["cells_step"]
        roof, hbm = rooflines(r, d, K, p_int32, hbm_peak, measured)
        cfg = config_dict(args.config, n)
        line = {
            "metric": METRIC, "value": cells_all * K / (dev_ms * 1e-3) / 1e9, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": args.warmup,
            "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u16x2 (score-safe pairs) / int32", "data": "synthetic",
            "config": cfg,
            "cells_per_gpu": cells_step,
            "cpu_binding": (f"each rank on the {numa} CPUs local to its GPU (NVML)" if numa else "none"),
            "alignments_per_s": pairs_all * K / (dev_ms * 1e-3),
            "wall_ms_per_step": wall_ms / K,
            "kernel_ms_per_step": {"fill": r["fill_ms"] / K, "traceback": r["trace_ms"] / K, "scan_compact": r["post_ms"] / K,
                                   "serial_total": r["serial_ms"] / K, "fill_launches": r["fill_launches"] // K,
                                   "note": "second pass, one chunk at a time, CUDA events around each phase on its stream"},
            "gpu_launches": int(r["launches"]),
            "clocks": clocks,
            "roofline": roof,
            "roofline_hbm": hbm,
            "int_peaks_gops": peaks,
        }
        if not args.no_e2e:
            line["e2e"] = {"value": cells_all * K / (e2e_ms * 1e-3) / 1e9, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
                           "d2h_bytes_per_step": int(d2h), "ms_per_step": e2e_ms / K,
                           "alignments_per_s": pairs_all * K / (e2e_ms * 1e-3),
                           "host_cpu_ns_per_pair": e2e_cpu_s / (n * K) * 1e9,
                           "host_cpu_note": "process CPU time (all threads of rank 0, spinning waits included) per pair inside the timed e2e calls"}
            line["e2e_pageable"] = {"value": cells_all * pg_steps / (pg_ms * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": pg_ms / pg_steps,
                                    "steps": pg_steps,
                                    "note": "the same call with the caller's arena and result arrays in ordinary pageable memory (what a JVM "
                                            "direct ByteBuffer is): every cudaMemcpyAsync is staged by the driver"}
            line["score_only_e2e"] = {"value": cells_all * K / (so_ms * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": so_ms / K,
                                      "d2h_bytes_per_step": 12 * n,
                                      "note": "mgl_sw_score_batch, host buffers in and out: the fill without checkpoints, start cell and score per pair"}

    # ---- sub-records (fixed 3 steps each, whatever --steps says)
    if not args.no_sub and args.config == 2 and not args.pairs:
        if world == 1:
            subs = {}
            for c in (3, 4, 5):
                rec = sub_config(sw, c, SUB_PAIRS[c], rank, dist, local, p_int32, hbm_peak, measured)
                subs[str(c)] = rec
            if line is not None:
                line["configs"] = subs
                line["configs_note"] = ("BASELINE.json configs 3, 4, 5 measured like the headline (device-resident value, per-kernel fill time, "
                                        "e2e through mgl_sw_align_batch), 3 steps after 3 warm-up steps each; config 4 at 4 000 of its 10 000 pairs, "
                                        "config 5 at 1 M pairs standing in for 100 M (1.5e13 cells ~ 3 GPU-hours at 1.4 TCUPS)")
        else:
            rec = sub_config(sw, 5, SUB_PAIRS[5], rank, dist, local, p_int32, hbm_peak, measured)
            if line is not None:
                line["config5"] = rec
                line["config5_note"] = ("BASELINE.json configs[4], the scaling sweep's workload: 1 M mixed pairs PER GPU (weak scaling) standing in "
                                        "for 100 M pairs in total, which would be ~3 GPU-hours at 1.4 TCUPS")
    if world > 1 and not args.no_sub and args.config == 2 and not args.pairs:
        sw.close()                                       # every rank gives its GPU back before one process drives them all
        barrier(dist, local)
        cpu_barrier(dist)
        if rank == 0:
            try:
                line["multi_handle"] = multi_handle_record(world, n)
            except Exception as ex:                      # the headline must not die with the sub-record
                line["multi_handle"] = {"error": repr(ex)}
        cpu_barrier(dist)
    if line is not None and world == 1 and not args.no_sub and args.config == 2 and not args.pairs:
        try:
            line["pairhmm"] = pairhmm_record(local, peaks.get("ffma_f32"))
        except Exception as ex:
            line["pairhmm"] = {"error": repr(ex)}
    if line is not None and world == 1 and not args.no_cpu_baseline:
        cb = cpu_baseline(d)
        if cb:
            line["cpu_baseline"] = cb
    if line is not None:
        emit((line))
    sw.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

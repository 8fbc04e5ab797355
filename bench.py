#!/usr/bin/env python
"""bench.py -- the headline benchmark: Smith-Waterman (fill + traceback + CIGAR)
throughput on BASELINE.json config 2 (1 M synthetic 150 bp reads vs 300-500 bp
haplotypes, scores (10,-15,-30,-5), SOFTCLIP) per GPU.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--pairs P] [--config C]

A step is one pass of the hot path over the whole batch.  `value` is GCUPS with
inputs resident in HBM (CUDA events on the launching stream, summed over the
batch's kernels, max over ranks); `e2e` is the same metric through the C-ABI
mgl_sw_align_batch with pinned HOST buffers in and out.  N > 1: one process per
GPU (torchrun), each rank aligns its own shard of the synthetic stream, no
data-path collective (pairs are independent) -> weak scaling.

--impl reference times the reference's own CPU implementation (oracle/_ref:
the unmodified align_avx / align_scalar behind the JNI dispatch rule, all host
threads) on a bounded sample of the same workload.
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
# DRAM bytes per DP cell of the fill kernel from the ncu --set full capture of the same command
# (profiles/r01m_fill_trace_full_config2.txt: K = 20 launch, dram read + write 3.756 GB for 3.365 Gcells)
NCU_FILL_DRAM_BYTES_PER_CELL = {2: 1.116}


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
    ap.add_argument("--no-numa-bind", action="store_true", help="N > 1: do not bind each rank to the CPUs local to its GPU")
    ap.add_argument("--chunk-pairs", type=int, default=0, help="engine chunk size (0: library default)")
    ap.add_argument("--pipe-wide", type=int, default=-1, help="long-pair kernel on eight warps per pair: 0 never, 1 auto, 2 always (-1: library default)")
    ap.add_argument("--pack", type=int, default=-1, help="2-bit transport of all-ACGT arena ranges: 0 never, 1 when the planner is ahead, 2 always (-1: library default)")
    ap.add_argument("--pack-lead", type=int, default=-1, help="chunks a planner must be ahead of the upload to pack (-1: library default)")
    ap.add_argument("--plan-threads", type=int, default=0, help="host planner threads (0: library default)")
    return ap.parse_args()


DEFAULT_PAIRS = {2: 1_000_000, 3: 100_000, 4: 1_000, 5: 1_000_000}
WORKLOAD = {2: "config 2: 150 bp reads vs 300-500 bp haplotypes, (10,-15,-30,-5), SOFTCLIP",
            3: "config 3: 300-1000 bp hap-to-ref, (200,-150,-260,-11), SOFTCLIP/INDEL",
            4: "config 4: 5-10 kbp pairs, (200,-150,-260,-11), SOFTCLIP",
            5: "config 5: mixed 90 % c2 / 9.9 % c3 / 0.1 % c4 shapes"}


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
        return rank, world, local, dist
    return rank, world, local, None


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
    threads = O.ref_max_threads()
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
    """--impl reference: the reference's CPU implementation, rank 0 only."""
    if rank != 0:
        return
    from mgl_b200 import synth
    from oracle import sw_oracle as O
    n = args.pairs or DEFAULT_PAIRS[args.config]
    d = synth.generate(args.config, n)
    if not O.have_ref():
        emit(({"impl": "reference", "unavailable": "oracle/_ref/libmgl_sw_ref.so was not built (no reference sources on this box)"}))
        return
    threads = O.ref_max_threads()
    probe = min(n, max(64, 40 * threads))
    secs, _ = cpu_reference_run(d, probe, threads)
    per_step_budget = 60.0 / max(1, args.steps + args.warmup)
    n_sample = int(min(n, max(probe, probe / max(secs, 1e-6) * min(per_step_budget, 10.0))))
    for _ in range(args.warmup):
        cpu_reference_run(d, n_sample, threads)
    tot_s, tot_cells = 0.0, 0
    for _ in range(args.steps):
        s, c = cpu_reference_run(d, n_sample, threads)
        tot_s += s
        tot_cells += c
    val = tot_cells / tot_s / 1e9
    sample = f"each step = first {n_sample} pairs of the workload ({tot_cells / args.steps / 1e9:.2f} Gcells)"
    emit(({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": tot_s / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        "config": {"workload": WORKLOAD[args.config], "pairs_per_step": n_sample},
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
    d = synth.generate(args.config, n, first=rank * n, pinned=True)       # this rank's shard of the stream
    sw = mgl_b200.MicrosoftSmithWaterman(local)
    sw._ensure()                                                          # no GPU / no extension -> raises
    if args.chunk_pairs:
        sw.set_option("chunk_pairs", args.chunk_pairs)
    if args.pipe_wide >= 0:
        sw.set_option("pipe_wide", args.pipe_wide)
    if args.plan_threads:
        sw.set_option("plan_threads", args.plan_threads)
    if args.pack >= 0:
        sw.set_option("pack_arena", args.pack)
    if args.pack_lead >= 0:
        sw.set_option("pack_lead", args.pack_lead)

    # integer issue rates (roofline denominator), measured on this box
    peaks = {}
    if rank == 0:
        L = native.lib()
        import ctypes as C
        for k in range(L.mgl_sw_int_peak_count()):
            ops, mhz = C.c_double(), C.c_double()
            if L.mgl_sw_int_peak(local, k, C.byref(ops), C.byref(mhz)) == 0:
                peaks[L.mgl_sw_int_peak_name(k).decode()] = ops.value / 1e9

    # ---- value: inputs resident in HBM, kernels timed with CUDA events on their streams.
    # Pass 1 (the reported value): the engine's production schedule -- chunks alternate between two work slots, so the
    # latency-bound traceback of chunk c overlaps the fill of chunk c + 1; one event pair brackets the whole batch.
    # Pass 2 (roofline of the dominant kernel): one chunk at a time, every phase bracketed by events on its stream.
    sw.resident_prepare(d["arena"], d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"])
    sw.set_option("overlap", 1)
    for _ in range(args.warmup):
        sw.resident_run()
    barrier(dist, local)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    dev_ms = 0.0
    launches = 0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        st = sw.resident_run()
        dev_ms += st["device_ms"]
        launches += st["fill_launches"] + st["other_launches"]
    barrier(dist, local)
    wall_ms = (time.perf_counter() - t0) * 1e3
    cells_step = st["cells"]
    all_s16 = st["pairs_s32"] == 0 and st["pairs_s16"] > 0    # every pair in u16x2 arithmetic: two cells per lane-op
    ckpt_bytes_step = st["trace_bytes"]                 # checkpoint bytes the fill launches of one step write (planner's count)
    sw.set_option("overlap", 0)
    fill_ms = trace_ms = post_ms = serial_ms = 0.0
    fill_launches = 0
    for k in range(args.warmup + args.steps):
        st = sw.resident_run()
        if k >= args.warmup:
            fill_ms += st["fill_ms"]; trace_ms += st["trace_ms"]; post_ms += st["post_ms"]; serial_ms += st["device_ms"]
            fill_launches += st["fill_launches"]
    sw.set_option("overlap", 1)
    sw.resident_drop()

    # ---- e2e: host buffers through the C-ABI, copies inside the timed region
    e2e_ms = 0.0
    h2d = d2h = 0
    if not args.no_e2e:
        # CIGAR element capacity: 8 per pair covers the short-read configs; long pairs carry ~1 element per 40 bases
        total_len = int(d["ref_len"].astype(np.int64).sum() + d["alt_len"].astype(np.int64).sum())
        out = (native.pinned_array(n, np.int32), native.pinned_array(n + 1, np.uint64),
               native.pinned_array(max(8 * n, total_len // 8), np.uint32))
        for _ in range(min(args.warmup, 2)):
            sw.align_arrays(d["arena"], d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"], out=out)
        barrier(dist, local)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            r = sw.align_arrays(d["arena"], d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"], out=out)
            h2d, d2h = r.stats["h2d_bytes"], r.stats["d2h_bytes"]
        e2e_ms = (time.perf_counter() - t0) * 1e3
        barrier(dist, local)

    # ---- score only: same host buffers through mgl_sw_score_batch, no checkpoints, no traceback
    so_ms = 0.0
    if not args.no_e2e:
        so_out = native.pinned_array(3 * n, np.int32).reshape(n, 3)
        sw.score_arrays(d["arena"], d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"], out=so_out)
        barrier(dist, local)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            sw.score_arrays(d["arena"], d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"], out=so_out)
        so_ms = (time.perf_counter() - t0) * 1e3
        barrier(dist, local)

    clocks = sampler.stop() if rank == 0 else None     # sampled over all three timed passes (value, per-kernel, e2e)
    (dev_ms, fill_ms, wall_ms, e2e_ms, so_ms), (cells_all, pairs_all) = reduce_max_sum(
        dist, local, [dev_ms, fill_ms, wall_ms, e2e_ms, so_ms], [float(cells_step), float(n)])

    if rank == 0:
        K = args.steps
        value = cells_all * K / (dev_ms * 1e-3) / 1e9
        fill_cups = cells_step * K / (fill_ms * 1e-3)                    # this GPU's fill kernels
        p_int32 = max(peaks.get("iadd3_s32", 0.0), peaks.get("imnmx_s32", 0.0)) or None
        # the denominator in the arithmetic the kernel runs: a u16x2 lane-op is two 16-bit cell-ops
        p_int = (2.0 * p_int32 if all_s16 else p_int32) if p_int32 else None
        measured = {}
        try:
            measured = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = measured.get("hbm_gbs", 6650.0)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": args.warmup,
            "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u16x2 (score-safe pairs) / int32", "data": "synthetic",
            "config": {"workload": WORKLOAD[args.config], "pairs_per_gpu": n, "cells_per_gpu": cells_step,
                       "l2": "inputs_exceed_l2 (arena + trace per step >> 126 MB)", "sharding": "independent pairs, no collective",
                       "cpu_binding": (f"each rank on the {numa} CPUs local to its GPU (NVML)" if numa else "none")},
            "alignments_per_s": pairs_all * K / (dev_ms * 1e-3),
            "wall_ms_per_step": wall_ms / K,
            "kernel_ms_per_step": {"fill": fill_ms / K, "traceback": trace_ms / K, "scan_compact": post_ms / K,
                                   "serial_total": serial_ms / K, "fill_launches": fill_launches // K,
                                   "note": "second pass, one chunk at a time, CUDA events around each phase on its stream"},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "int", "achieved": fill_cups * ALGO_OPS_PER_CELL / 1e9, "peak": p_int, "unit": "Gop/s",
                         "frac": (fill_cups * ALGO_OPS_PER_CELL / 1e9 / p_int) if p_int else None,
                         "traffic": (NCU_FILL_DRAM_BYTES_PER_CELL[args.config] * cells_step / max(1, fill_launches // K))
                         if args.config in NCU_FILL_DRAM_BYTES_PER_CELL else None,
                         "note": "dominant kernel = sw_fill_kernel (all its launches of a step, timed with CUDA events on their stream in the "
                                 "one-chunk-at-a-time pass); algorithmic 10 int ops/cell (SURVEY.md 8d) x cells; peak = max(iadd3, imnmx) measured "
                                 "live by int_peak.cu (not in MEASURED_PEAKS.json)"
                                 + (", x 2 because every pair of this workload runs in u16x2 arithmetic (one lane-op = two 16-bit cell-ops); "
                                    "against the plain 32-bit peak (peak_s32) the fraction would be twice this" if all_s16 else
                                    "; pairs that run in u16x2 arithmetic do two cells per lane-op, so a mixed workload can exceed its 32-bit fraction")
                                 + "; traffic = ncu dram bytes per cell x cells per average fill launch",
                         "peak_s32": p_int32},
            "roofline_hbm": {"bound": "hbm", "achieved": ckpt_bytes_step * K / (fill_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                             "frac": ckpt_bytes_step * K / (fill_ms * 1e-3) / 1e9 / hbm_peak,
                             "algorithmic_bytes_per_cell": ckpt_bytes_step / cells_step,
                             "note": "checkpoint writes of the fill kernel (the only O(cells) HBM traffic on the path); peak = MEASURED_PEAKS.json hbm_gbs"
                                     + ("" if measured else " (fallback 6650)")},
            "int_peaks_gops": peaks,
        }
        if not args.no_e2e:
            line["e2e"] = {"value": cells_all * K / (e2e_ms * 1e-3) / 1e9, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
                           "d2h_bytes_per_step": int(d2h), "ms_per_step": e2e_ms / K,
                           "alignments_per_s": pairs_all * K / (e2e_ms * 1e-3)}
            line["score_only_e2e"] = {"value": cells_all * K / (so_ms * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": so_ms / K,
                                      "d2h_bytes_per_step": 12 * n,
                                      "note": "mgl_sw_score_batch, host buffers in and out: the fill without checkpoints, start cell and score per pair"}
        if world == 1 and not args.no_cpu_baseline:
            cb = cpu_baseline(d)
            if cb:
                line["cpu_baseline"] = cb
        emit((line))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

/*
 * sw_trace_kernel.cu -- the traceback's kernels: start cells (a lane group per pair; gap-free paths end there), the walks
 * of the other pairs (Walker, sw_trace.h: short pairs one rebuilt block per launch, the rest one lane per pair to the
 * end), and the small kernels that turn per-pair CIGAR scratch into one dense array or into text.
 * Replaces calculateCigar / calculateCigar_avx (sw.cpp:149-255, sw_avx.cpp:324-430).
 */
#include "sw_kernels.h"
#include "sw_trace.h"
#include "sw_config.h"

namespace mglsw {

/*
 * The start cell of one pair (select_start, sw_trace.h: sw.cpp:116-127 and :155-170) by a group of 8 lanes: lane g reads
 * the words of lane slots g, g + 8, ... of the captured last row -- several in flight, where one thread per pair walked
 * the whole row one dependent load at a time.  The scan keeps the FIRST of the candidates with the best (score, distance to the
 * diagonal) in its order -- last-column maximum first, then the last row by ascending column (it replaces only on a
 * strict improvement) -- so the lanes keep (score, distance, column) and the reduction prefers the smaller column among
 * equals; the last-column maximum wins every tie.  The same value in all 8 lanes.
 */
__device__ __forceinline__ StartCell start_cell_group(const UnitDesc &u, u32 half, bool s32, u32 G, u32 K, const u32 *colbest4,
                                                      const u32 *lastrow, int strategy, u32 gl, unsigned gmask)
{
    const int tl = (int)u.tl[half], ql = (int)u.ql;
    const u32 nslots = u.nstrips * G;
    const u32 *row = lastrow + u.lastrow_off + (size_t)half * nslots * K;
    const u32 sh = s32 ? 0u : 16u * half, msk = s32 ? 0xffffffffu : 0xffffu;
    const int mqe = (int)__ldcg(colbest4 + 2 * half), mqe_t = (int)__ldcg(colbest4 + 2 * half + 1);
    StartCell sc;
    if (strategy == OS_INDEL) {
        u32 slot, reg;
        col_to_slot(u, (u32)ql - 1u, slot, reg);
        sc.i = tl; sc.j = ql; sc.seg = 0; sc.score = (int)((__ldcg(row + (size_t)slot * K + reg) >> sh) & msk);
    } else if (strategy == OS_LEADING_INDEL) {
        sc.i = mqe_t; sc.j = ql; sc.seg = 0; sc.score = mqe;
    } else {
        /* lane g takes lane slots g, g + 8, ...: K consecutive words each, columns ascending; the 8 lanes of a group read
           neighbouring 4 K-byte runs of one row, so the sectors are shared through L1 (the row was written by an earlier launch) */
        /* first only the row's maximum (three instructions per word): a last row that stays below the last-column maximum --
           a read that ends inside its haplotype -- cannot be where the walk starts, and the scan with the tie rules is skipped */
        int rowmax = (int)0x80000000;
        for (u32 slot = gl; slot < nslots; slot += 8u) {
            const u32 cols = slot_cols(u, slot);
            const u32 *wp = row + (size_t)slot * K;
#pragma unroll 4
            for (u32 reg = 0; reg < cols; reg++) { const int s = (int)((wp[reg] >> sh) & msk); rowmax = s > rowmax ? s : rowmax; }
        }
#pragma unroll
        for (int m = 1; m < 8; m <<= 1) { const int o = __shfl_xor_sync(gmask, rowmax, m); rowmax = o > rowmax ? o : rowmax; }
        if (rowmax < mqe) { sc.i = mqe_t; sc.j = ql; sc.seg = 0; sc.score = mqe; return sc; }
        int best = (int)0x80000000, bd = 0x7fffffff, bj = 0x7fffffff;
        for (u32 slot = gl; slot < nslots; slot += 8u) {
            const u32 cols = slot_cols(u, slot);
            int j = (int)slot_first_col(u, slot) + 1;
            const u32 *wp = row + (size_t)slot * K;
#pragma unroll 4
            for (u32 reg = 0; reg < cols; reg++, j++) {
                const int s = (int)((wp[reg] >> sh) & msk);
                const int da = tl - j < 0 ? j - tl : tl - j;
                if (s > best || (s == best && da < bd)) { best = s; bd = da; bj = j; }
            }
        }
#pragma unroll
        for (int m = 1; m < 8; m <<= 1) {
            const int os = __shfl_xor_sync(gmask, best, m), od = __shfl_xor_sync(gmask, bd, m), oj = __shfl_xor_sync(gmask, bj, m);
            if (os > best || (os == best && (od < bd || (od == bd && oj < bj)))) { best = os; bd = od; bj = oj; }
        }
        const int db0 = mqe_t - ql < 0 ? ql - mqe_t : mqe_t - ql;
        if (best > mqe || (best == mqe && bd < db0)) { sc.i = tl; sc.j = bj; sc.seg = ql - bj; sc.score = best; }
        else { sc.i = mqe_t; sc.j = ql; sc.seg = 0; sc.score = mqe; }
    }
    return sc;
}

/*
 * 8 lanes per pair slot (2 v + h = half h of unit v).  SCORE_ONLY: the start cell and its H are the result
 * (mgl_sw_score_batch).  Otherwise the group also counts the mismatches along the diagonal from the start cell to row 0 /
 * column 0: if the scores say that the whole path is that diagonal (clean_path) lane 0 writes the pair's CIGAR and
 * offset and the pair is done -- no checkpoint is read; else the pair joins the worklist of sw_traceback_kernel.
 */
template <bool SCORE_ONLY>
__global__ void __launch_bounds__(128) sw_start_kernel(const __grid_constant__ TraceArgs a)
{
    const u32 sid = (blockIdx.x * 128u + threadIdx.x) >> 3;
    const u32 gl = threadIdx.x & 7u;
    const unsigned gmask = 0xffu << (threadIdx.x & 24u);
    bool push = false, is_long = false;
    StartCell sc;
    sc.i = sc.j = sc.seg = sc.score = 0;
    if (sid < 2u * a.n_units) {
        const UnitDesc &u = a.units[sid >> 1];
        const u32 half = sid & 1u;
        const u32 p = u.pair[half];
        if ((u.flags & UNIT_ACTIVE) && p != NO_PAIR) {
            const bool s32 = (u.flags & UNIT_S32) != 0;
            const u32 G = (u.flags & UNIT_G32) ? MGLSW_G_LONG : MGLSW_G_SHORT;
            is_long = (u.flags & UNIT_G32) != 0 || a.no_rounds != 0u;
            const CellConsts &c = s32 ? a.c32 : a.c16;
            const int strategy = a.strategy[p];
            sc = start_cell_group(u, half, s32, G, u.regs, a.colbest + (size_t)(sid >> 1) * 4u, a.lastrow, strategy, gl, gmask);
            if (SCORE_ONLY) {
                if (gl == 0) {
                    a.score3[3 * (size_t)p] = sc.score - c.bias - c.s_match;
                    a.score3[3 * (size_t)p + 1] = sc.i;
                    a.score3[3 * (size_t)p + 2] = sc.j;
                }
            } else {
                const int indel = (u.flags & UNIT_INDEL) ? 1 : 0;
                if (!clean_path_possible(sc, c, indel)) push = gl == 0;
                else {
                    const int L = sc.i < sc.j ? sc.i : sc.j;
                    const uint8_t *rf = a.arena + u.ref_off[half] + (sc.i - L), *al = a.arena + u.alt_off[half] + (sc.j - L);
                    int nm = 0;
                    for (int t = (int)gl; t < L; t += 8) nm += __ldg(rf + t) != __ldg(al + t) ? 1 : 0;
#pragma unroll
                    for (int m = 1; m < 8; m <<= 1) nm += __shfl_xor_sync(gmask, nm, m);
                    if (gl == 0) {
                        PairResult r;
                        if (clean_path(sc, strategy, c, indel, nm, a.scratch + a.ops_end[p], r)) {
                            a.offsets[p] = r.offset;
                            a.n_ops[p] = r.n_ops;
                        } else push = true;
                    }
                }
            }
        }
    }
    if (SCORE_ONLY) return;
    /*
     * Two lists in one array: pairs of lane groups of 8 from the front (they go through the round kernels), pairs of
     * whole-warp units from the back (sw_traceback_kernel walks them from the start cell).  One atomic per warp and
     * list: its (at most four) pairs take consecutive entries.
     */
    const u32 wl = threadIdx.x & 31u;
#pragma unroll
    for (int which = 0; which < 2; which++) {
        const unsigned pm = __ballot_sync(0xffffffffu, push && (int)is_long == which);
        if (pm == 0u) continue;
        u32 base = 0;
        if (wl == (u32)(__ffs(pm) - 1)) base = atomicAdd(a.work_count + (which ? MGLSW_WC_LONG : 0), (u32)__popc(pm));
        base = __shfl_sync(0xffffffffu, base, __ffs(pm) - 1);
        if (push && (int)is_long == which) {
            u32 k = base + (u32)__popc(pm & ((1u << wl) - 1u));
            if (which) k = 2u * a.n_units - 1u - k;
            a.worklist[k] = sid;
            a.startcell[k] = make_int4(sc.i, sc.j, sc.seg, sc.score);
        }
    }
}

#ifndef MGLSW_TRACE_THREADS
#define MGLSW_TRACE_THREADS 128
#endif
#ifndef MGLSW_TRACE_MINB
#define MGLSW_TRACE_MINB (512 / MGLSW_TRACE_THREADS)    /* 128 registers per thread */
#endif

/* one walk's identity from its pair slot, and the Walker set up on it: from the start cell, or from a saved state */
struct WalkCtx { u32 p, flags, strips; };
__device__ __forceinline__ WalkCtx walk_open(const TraceArgs &a, u32 sid, const int4 *start, const u32 *saved, Walker &w, u64 *nib)
{
    const UnitDesc &u = a.units[sid >> 1];
    const u32 half = sid & 1u;
    WalkCtx x;
    x.p = u.pair[half]; x.flags = u.flags; x.strips = u.nstrips;
    const bool s32 = (u.flags & UNIT_S32) != 0;
    const u32 G = (u.flags & UNIT_G32) ? MGLSW_G_LONG : MGLSW_G_SHORT;
    u32 *end = a.scratch + a.ops_end[x.p];
    if (start) {
        StartCell sc;
        sc.i = start->x; sc.j = start->y; sc.seg = start->z; sc.score = start->w;
        w.begin(u, half, s32, G, u.regs, a.ckpt, a.arena, s32 ? a.c32 : a.c16, sc, a.strategy[x.p], end, nib, (u32)MGLSW_TRACE_THREADS);
    } else {
        w.resume(u, half, s32, G, u.regs, a.ckpt, a.arena, s32 ? a.c32 : a.c16, saved, a.strategy[x.p], end, nib, (u32)MGLSW_TRACE_THREADS);
    }
    return x;
}

/*
 * Round r of the pairs of lane groups of 8 (short paths: a few hops between two rebuilt blocks).  One lane per entry
 * of round r's list; every entry needs exactly one block rebuilt now, so the replay site -- two thirds of the
 * traceback's instructions -- runs with every live lane, where one launch that kept a pair on its lane until the end
 * ran it with a third of them (the pairs of a warp need 1 to 10 rebuilds, 2.7 on average, and the warp waited for the
 * slowest).  Then the walk hops on; a pair that reaches the matrix edge is written out, the others are put away
 * (WalkState) on round r + 1's list, densely.  FIRST: entries come from sw_start_kernel (start cells); the walk
 * begins with hops.
 */
template <bool FIRST>
__global__ void __launch_bounds__(MGLSW_TRACE_THREADS, MGLSW_TRACE_MINB) sw_trace_round_kernel(const __grid_constant__ TraceArgs a, const u32 round)
{
    __shared__ u64 s_nib[MGLSW_TRACE_BUF * MGLSW_TRACE_THREADS];
    const u32 k = blockIdx.x * blockDim.x + threadIdx.x;
    const u32 count = __ldcg(a.work_count + round);
    if ((k & ~31u) >= count) return;                /* whole warps leave; a lane past the end stays with its warp, idle */
    const bool live = k < count;
    const uint4 *in = a.wstate[round & 1u];
    uint4 *out = a.wstate[(round + 1u) & 1u];
    Walker w;
    w.done = true;
    WalkCtx x;
    x.p = 0; x.flags = 0; x.strips = 0;
    u32 sid = 0;
    if (live) {
        if (FIRST) {
            sid = a.worklist[k];
            const int4 s4 = a.startcell[k];
            x = walk_open(a, sid, &s4, nullptr, w, s_nib + threadIdx.x);
            w.hop_phase<MGLSW_HOP_SHORT, false>();
        } else {
            __align__(16) u32 rec[12];
            *reinterpret_cast<uint4 *>(rec) = __ldcg(in + 3 * (size_t)k);
            *reinterpret_cast<uint4 *>(rec + 4) = __ldcg(in + 3 * (size_t)k + 1);
            *reinterpret_cast<uint4 *>(rec + 8) = __ldcg(in + 3 * (size_t)k + 2);
            sid = rec[0];
            x = walk_open(a, sid, nullptr, rec, w, s_nib + threadIdx.x);
        }
    }
    __syncwarp();
    if (!w.done) { w.build_phase(); w.hop_phase<MGLSW_HOP_SHORT, false>(); }
    if (live && w.done) {
        const PairResult r = w.finish();
        a.offsets[x.p] = r.offset;
        a.n_ops[x.p] = r.n_ops;
    }
    const bool push = live && !w.done;
    const unsigned pm = __ballot_sync(0xffffffffu, push);
    if (pm == 0u) return;
    const u32 wl = threadIdx.x & 31u;
    u32 base = 0;
    if (wl == (u32)(__ffs(pm) - 1)) base = atomicAdd(a.work_count + round + 1u, (u32)__popc(pm));
    base = __shfl_sync(0xffffffffu, base, __ffs(pm) - 1);
    if (push) {
        const u32 q = base + (u32)__popc(pm & ((1u << wl) - 1u));
        __align__(16) u32 rec[12];
        w.save(sid, rec);
        __stcg(out + 3 * (size_t)q, *reinterpret_cast<uint4 *>(rec));
        __stcg(out + 3 * (size_t)q + 1, *reinterpret_cast<uint4 *>(rec + 4));
        __stcg(out + 3 * (size_t)q + 2, *reinterpret_cast<uint4 *>(rec + 8));
    }
}

/*
 * One lane per pair until its walk ends: the pairs the round kernels have not finished after MGLSW_TRACE_ROUNDS
 * rebuilds (RESUME: from their saved state; a few per cent of config 2's), and the pairs of whole-warp units (from their
 * start cells, at the back of the worklist).  Each lane drives a Walker; two choices per warp, so that its lanes keep
 * running the same code [B200, profiles/r22b, r22c; traceback alone, before the round kernels]:
 * - lane groups of 8 meet at a vote once per round, so that the lanes that need a block rebuilt run the ONE replay site
 *   together.  Left to the hardware's reconvergence the lanes drifted apart: the replay ran 19 times per warp with 4
 *   lanes each where the slowest pair needs 8 (config 2: 5.6 -> 3.8 ms per 1 M pairs; profiles/r19a).  On long paths
 *   every round lasts as long as the longest clean stretch among 32 lanes, so warps of whole-warp units let their
 *   lanes run free, as the hardware reconverges them (config 3: 2.9 ms per 100 k pairs free, 5.1 ms met);
 * - hops of MGLSW_HOP_BLOCKS blocks per round trip pay where the chain of round trips is what a pair waits for
 *   (config 4's shapes; config 5 33 -> 29 ms per 1 M pairs); elsewhere their instructions cost more than the round
 *   trips they save (config 3 2.9 -> 4.8 ms; config 2's overlapped step 15.8 -> 16.2 ms).
 * (Lanes that take the next entry off the list when their pair is done -- a grid of a third or a quarter of the list --
 * keep the replay site fuller but were slower, alone and in the overlapped step: profiles/r21a.)
 */
/* HAS_WIDE false: no pair of MGLSW_HOP_WIDE_MIN_STRIPS strips in the chunk, the kernel carries no four-block hop; SMALL: a small
   chunk's one launch, whose list holds lane groups of 8 as well (otherwise the back list is whole-warp units only, and the
   kernel carries neither the meeting point nor the two-block hop: its lanes run apart, so every instruction it does not
   hold is instruction cache they do not fight over) */
template <bool HAS_WIDE, bool RESUME, bool SMALL>
__global__ void __launch_bounds__(MGLSW_TRACE_THREADS, MGLSW_TRACE_MINB) sw_traceback_kernel(const __grid_constant__ TraceArgs a)
{
    __shared__ u64 s_nib[MGLSW_TRACE_BUF * MGLSW_TRACE_THREADS];   /* the replayed block: records and boundary H values, one column per thread */
    const u32 k = blockIdx.x * blockDim.x + threadIdx.x;
    const u32 count = __ldcg(a.work_count + (RESUME ? MGLSW_TRACE_ROUNDS : MGLSW_WC_LONG));
    if ((k & ~31u) >= count) return;                /* whole warps leave; a lane past the end stays with its warp, idle */
    const bool live = k < count;
    Walker w;
    w.done = true;
    WalkCtx x;
    x.p = 0; x.flags = 0; x.strips = 0;
    if (live) {
        if (RESUME) {
            const uint4 *in = a.wstate[MGLSW_TRACE_ROUNDS & 1];
            __align__(16) u32 rec[12];
            *reinterpret_cast<uint4 *>(rec) = __ldcg(in + 3 * (size_t)k);
            *reinterpret_cast<uint4 *>(rec + 4) = __ldcg(in + 3 * (size_t)k + 1);
            *reinterpret_cast<uint4 *>(rec + 8) = __ldcg(in + 3 * (size_t)k + 2);
            x = walk_open(a, rec[0], nullptr, rec, w, s_nib + threadIdx.x);
        } else {
            const u32 e = 2u * a.n_units - 1u - k;
            const int4 s4 = a.startcell[e];
            x = walk_open(a, a.worklist[e], &s4, nullptr, w, s_nib + threadIdx.x);
        }
    }
    const bool meet = RESUME || (SMALL && __all_sync(0xffffffffu, !live || !(x.flags & UNIT_G32)));
    const bool wide = HAS_WIDE && __all_sync(0xffffffffu, !live || x.strips >= MGLSW_HOP_WIDE_MIN_STRIPS);
    for (;;) {
        if (!w.done) {                                                 /* each lane on its own */
            if (RESUME) w.hop_phase<MGLSW_HOP_SHORT, false>();         /* the round kernels' leftovers: lane groups of 8 */
            else if (wide) w.hop_phase<MGLSW_HOP_BLOCKS>();
            else if (SMALL && meet) w.hop_phase<MGLSW_HOP_SHORT>();    /* a small chunk's lane groups of 8 */
            else w.hop_phase<1>();
        }
        if (meet) { if (__ballot_sync(0xffffffffu, !w.done) == 0u) break; }   /* leaves when every lane is done */
        else if (w.done) break;
        if (!w.done) w.build_phase();                                  /* met: the lanes that need it, together */
    }
    if (!live) return;
    const PairResult r = w.finish();
    a.offsets[x.p] = r.offset;
    a.n_ops[x.p] = r.n_ops;
}

/*
 * The exact-substring shortcut, opt-in.  The reference skips it on purpose -- "we skip substring
 * check since it performed by SWNativeAlignerWrapper" (...MicrosoftSmithWaterman.cpp:60) -- because GATK's wrapper
 * around the native binding (external to the reference: org.broadinstitute.hellbender...SWNativeAlignerWrapper.align)
 * answers SOFTCLIP and IGNORE calls itself when the alt occurs verbatim in the ref: offset = the LAST index at which
 * it occurs, CIGAR = <alt length>M.  A batched entry that bypasses that wrapper has no such pre-check.  (On every case
 * tried the DP's own tie rules -- the last-column maximum with ">=", sw.cpp:100-104 -- pick the last occurrence as
 * well, so the two agree; the option exists so that the wrapper's rule, not an observation, decides.)  With the option on,
 * this kernel overrides the traceback's result for exactly those pairs: one warp per pair, 32 candidate starts at a
 * time from the back of the ref, so the first group with a hit holds the last occurrence.
 */
__global__ void __launch_bounds__(128) sw_exact_match_kernel(const __grid_constant__ TraceArgs a)
{
    const u32 wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (wid >= 2u * a.n_units) return;
    const UnitDesc &u = a.units[wid >> 1];
    const u32 half = wid & 1u;
    const u32 p = u.pair[half];
    if (!(u.flags & UNIT_ACTIVE) || p == NO_PAIR || (half == 1u && (u.flags & UNIT_S32))) return;
    const int strategy = a.strategy[p];
    if (strategy != OS_SOFTCLIP && strategy != OS_IGNORE) return;
    const int tl = (int)u.tl[half], ql = (int)u.ql;
    if (ql > tl) return;
    const uint8_t *ref = a.arena + u.ref_off[half], *alt = a.arena + u.alt_off[half];
    for (int base = tl - ql; base >= 0; base -= 32) {
        const int i = base - (int)lane;
        bool hit = i >= 0;
        if (hit) {
            const uint8_t *r = ref + i;
            for (int k = 0; k < ql; k++) if (__ldg(r + k) != __ldg(alt + k)) { hit = false; break; }
        }
        const u32 m = __ballot_sync(0xffffffffu, hit);
        if (m) {
            if (lane == 0) {
                const int idx = base - (__ffs(m) - 1);                   /* the lowest lane with a hit = the largest start */
                a.offsets[p] = idx;
                a.n_ops[p] = 1u;
                a.scratch[a.ops_end[p] - 1u] = ((u32)ql << 4) | OP_M;
            }
            return;
        }
    }
}

cudaError_t launch_exact_match(const TraceArgs &a, cudaStream_t stream)
{
    if (a.n_units == 0) return cudaSuccess;
    sw_exact_match_kernel<<<(2u * a.n_units * 32u + 127) / 128, 128, 0, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_endpoints(const TraceArgs &a, cudaStream_t stream)
{
    if (a.n_units == 0) return cudaSuccess;
    sw_start_kernel<true><<<(2u * a.n_units * 8u + 127) / 128, 128, 0, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_start(const TraceArgs &a, cudaStream_t stream)
{
    if (a.n_units == 0) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(a.work_count, 0, MGLSW_WC_WORDS * sizeof(u32), stream);
    if (e != cudaSuccess) return e;
    sw_start_kernel<false><<<(2u * a.n_units * 8u + 127) / 128, 128, 0, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_traceback(const TraceArgs &a, bool has_short, bool has_long, bool has_wide, cudaStream_t stream)
{
    if (a.n_units == 0) return cudaSuccess;
    /* the lists' lengths are on the device: every grid covers every pair slot, warps beyond a list leave at once */
    const u32 blocks = (2u * a.n_units + MGLSW_TRACE_THREADS - 1) / MGLSW_TRACE_THREADS;
    if (has_short) {
        sw_trace_round_kernel<true><<<blocks, MGLSW_TRACE_THREADS, 0, stream>>>(a, 0u);
        for (u32 r = 1; r < MGLSW_TRACE_ROUNDS; r++) sw_trace_round_kernel<false><<<blocks, MGLSW_TRACE_THREADS, 0, stream>>>(a, r);
        sw_traceback_kernel<false, true, false><<<blocks, MGLSW_TRACE_THREADS, 0, stream>>>(a);
    }
    if (has_long) {
        /* whole-warp units hold one pair each; a small chunk has every pair slot on this list */
        const u32 lb = ((a.no_rounds ? 2u : 1u) * a.n_units + MGLSW_TRACE_THREADS - 1) / MGLSW_TRACE_THREADS;
        if (a.no_rounds) {
            if (has_wide) sw_traceback_kernel<true, false, true><<<lb, MGLSW_TRACE_THREADS, 0, stream>>>(a);
            else sw_traceback_kernel<false, false, true><<<lb, MGLSW_TRACE_THREADS, 0, stream>>>(a);
        } else if (has_wide) sw_traceback_kernel<true, false, false><<<lb, MGLSW_TRACE_THREADS, 0, stream>>>(a);
        else sw_traceback_kernel<false, false, false><<<lb, MGLSW_TRACE_THREADS, 0, stream>>>(a);
    }
    return cudaGetLastError();
}

/*
 * Exclusive scan of the per-pair element counts, two levels.  Level 1: each block scans 1024
 * counts (coalesced 16-byte loads, warp shuffles) into block-local offsets and one block total.
 * Level 2: one block scans the block totals in place and writes the grand total to starts[n].
 * The compaction adds the two.
 */
#define SCAN_TILE 1024u
__global__ void __launch_bounds__(256) sw_scan_tiles_kernel(const u32 *__restrict__ n_ops, u32 *__restrict__ starts,
                                                            u32 *__restrict__ tile_sum, u32 n)
{
    __shared__ u32 wsum[8];
    const u32 base = blockIdx.x * SCAN_TILE + threadIdx.x * 4u;
    u32 v[4];
#pragma unroll
    for (int k = 0; k < 4; k++) v[k] = base + k < n ? n_ops[base + k] : 0u;
    const u32 mine = v[0] + v[1] + v[2] + v[3];
    u32 incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const u32 x = __shfl_up_sync(0xffffffffu, incl, d); if ((threadIdx.x & 31u) >= (u32)d) incl += x; }
    if ((threadIdx.x & 31u) == 31u) wsum[threadIdx.x >> 5] = incl;
    __syncthreads();
    u32 woff = 0, total = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) { if ((u32)w < (threadIdx.x >> 5)) woff += wsum[w]; total += wsum[w]; }
    u32 run = woff + incl - mine;
#pragma unroll
    for (int k = 0; k < 4; k++) { if (base + k < n) starts[base + k] = run; run += v[k]; }
    if (threadIdx.x == 0) tile_sum[blockIdx.x] = total;
}

__global__ void __launch_bounds__(1024) sw_scan_totals_kernel(u32 *__restrict__ tile_sum, u32 *__restrict__ starts, u32 n_tiles, u32 n)
{
    __shared__ u32 part[1024];
    const u32 tid = threadIdx.x;
    const u32 per = (n_tiles + 1023u) / 1024u;
    const u32 lo = tid * per < n_tiles ? tid * per : n_tiles;
    const u32 hi = lo + per < n_tiles ? lo + per : n_tiles;
    u32 sum = 0;
    for (u32 k = lo; k < hi; k++) sum += tile_sum[k];
    part[tid] = sum;
    __syncthreads();
    for (u32 d = 1; d < 1024; d <<= 1) {
        const u32 v = tid >= d ? part[tid - d] : 0u;
        __syncthreads();
        part[tid] += v;
        __syncthreads();
    }
    u32 run = part[tid] - sum;
    for (u32 k = lo; k < hi; k++) { const u32 c = tile_sum[k]; tile_sum[k] = run; run += c; }
    if (tid == 1023) starts[n] = part[1023];
}

cudaError_t launch_scan_counts(const u32 *n_ops, u32 *starts, u32 *tile_sum, u32 n, cudaStream_t stream)
{
    const u32 n_tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    if (n_tiles) sw_scan_tiles_kernel<<<n_tiles, 256, 0, stream>>>(n_ops, starts, tile_sum, n);
    sw_scan_totals_kernel<<<1, 1024, 0, stream>>>(tile_sum, starts, n_tiles, n);
    return cudaGetLastError();
}

__global__ void __launch_bounds__(256) sw_compact_kernel(const u32 *__restrict__ scratch, const u64 *__restrict__ ops_end,
                                                         const u32 *__restrict__ n_ops, const u32 *__restrict__ starts,
                                                         const u32 *__restrict__ tile_off, u32 *__restrict__ ops_out, u32 n)
{
    const u32 p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const u32 cnt = n_ops[p];
    const u32 *src = scratch + (ops_end[p] - cnt);
    u32 *dst = ops_out + tile_off[p / SCAN_TILE] + starts[p];
    for (u32 k = 0; k < cnt; k++) dst[k] = src[k];
}

cudaError_t launch_compact(const u32 *scratch, const u64 *ops_end, const u32 *n_ops, const u32 *starts, const u32 *tile_off,
                           u32 *ops_out, u32 n, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    sw_compact_kernel<<<(n + 255) / 256, 256, 0, stream>>>(scratch, ops_end, n_ops, starts, tile_off, ops_out, n);
    return cudaGetLastError();
}

/*
 * Text CIGARs on the device: what the reference builds with to_string(length) + state per
 * element, elements of length <= 0 skipped (sw.cpp:251-252; the traceback never emits those).  Two passes over a
 * pair's packed elements: count the characters, then (after the same scan that places packed elements) write them.
 */
__device__ __forceinline__ u32 dec_digits(u32 v)
{
    u32 d = 1;
    while (v >= 10u) { v /= 10u; d++; }
    return d;
}

__global__ void __launch_bounds__(256) sw_text_len_kernel(const u32 *__restrict__ scratch, const u64 *__restrict__ ops_end,
                                                          const u32 *__restrict__ n_ops, u32 *__restrict__ n_chars, u32 n)
{
    const u32 p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const u32 cnt = n_ops[p];
    const u32 *src = scratch + (ops_end[p] - cnt);
    u32 chars = 0;
    for (u32 k = 0; k < cnt; k++) chars += dec_digits(src[k] >> 4) + 1u;
    n_chars[p] = chars;
}

__global__ void __launch_bounds__(256) sw_text_format_kernel(const u32 *__restrict__ scratch, const u64 *__restrict__ ops_end,
                                                             const u32 *__restrict__ n_ops, const u32 *__restrict__ starts,
                                                             const u32 *__restrict__ tile_off, char *__restrict__ text_out, u32 n)
{
    const u32 p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const u32 cnt = n_ops[p];
    const u32 *src = scratch + (ops_end[p] - cnt);
    char *dst = text_out + tile_off[p / SCAN_TILE] + starts[p];
    for (u32 k = 0; k < cnt; k++) {
        const u32 e = src[k];
        u32 len = e >> 4;
        const u32 d = dec_digits(len);
        for (u32 q = d; q-- > 0;) { dst[q] = (char)('0' + len % 10u); len /= 10u; }
        dst[d] = "MIDNSHP=X"[e & 15u];                        /* BAM numbering, as mgl_sw_format_cigar */
        dst += d + 1u;
    }
}

cudaError_t launch_text_len(const u32 *scratch, const u64 *ops_end, const u32 *n_ops, u32 *n_chars, u32 n, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    sw_text_len_kernel<<<(n + 255) / 256, 256, 0, stream>>>(scratch, ops_end, n_ops, n_chars, n);
    return cudaGetLastError();
}

cudaError_t launch_text_format(const u32 *scratch, const u64 *ops_end, const u32 *n_ops, const u32 *starts, const u32 *tile_off,
                               char *text_out, u32 n, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    sw_text_format_kernel<<<(n + 255) / 256, 256, 0, stream>>>(scratch, ops_end, n_ops, starts, tile_off, text_out, n);
    return cudaGetLastError();
}

}  // namespace mglsw

/*
 * sw_kernels.h -- host-callable launchers of every CUDA kernel on the path.
 * Implemented in sw_fill_inst_*.cu / sw_fill_dispatch.cu / sw_trace_kernel.cu /
 * int_peak.cu; called by the engine (sw_engine.cu).
 */
#ifndef MGL_B200_SW_KERNELS_H
#define MGL_B200_SW_KERNELS_H
#include <cuda_runtime.h>
#include "sw_fill.h"

namespace mglsw {

/* fill: fails with cudaErrorInvalidValue when no kernel was built for (kind, G, K) */
cudaError_t launch_fill_any(int kind, int G, int K, int pipe, const FillArgs &a, u32 first_unit, cudaStream_t stream);

struct TraceArgs {
    const UnitDesc *units;
    const u64 *ops_end;       /* per pair: end of its scratch region (words)    */
    const int *strategy;      /* per pair                                       */
    const u32 *ckpt;          /* checkpoints written by the fill                */
    const uint8_t *arena;     /* sequences (padded device arena)                */
    CellConsts c16, c32;      /* scores (+ bias) of the two arithmetics         */
    const u32 *colbest;       /* per unit 4 words, written by the fill          */
    const u32 *lastrow;
    u32 *scratch;             /* CIGAR elements, written downwards from ops_end */
    int *offsets;             /* out, per pair                                  */
    u32 *n_ops;               /* out, per pair                                  */
    int *score3;              /* out, score-only runs: per pair {score, end_ref, end_alt} */
    u32 n_units;              /* units of the chunk (padding units included)    */
    /*
     * launch_start -> launch_traceback: the pairs whose path has a gap.  worklist / startcell hold two lists: entries
     * 0 .. work_count[0] - 1 the pairs of lane groups of 8, entries 2 n_units - 1 downwards (work_count[MGLSW_WC_LONG] of
     * them) the pairs of whole-warp units.  work_count[r] = length of round r's list (sw_trace_round_kernel), whose
     * entries are walks put away in wstate[r & 1] (three 16-byte words each: WalkState, sw_trace.h).
     */
    u32 *work_count;          /* MGLSW_WC_WORDS words, zeroed by launch_start    */
    u32 *worklist;            /* 2 * n_units: pair slot (2 * unit + half)        */
    int4 *startcell;          /* 2 * n_units: {i, j, seg, score} (StartCell, sw_trace.h) */
    uint4 *wstate[2];         /* 2 * n_units * 3 each                            */
    u32 no_rounds;            /* a small chunk (one pair of alignNative, say): every pair goes on the back list and is walked by one
                                 launch -- six more launches would cost more than the lanes they save */
};
/*
 * The traceback's start cell of every pair (sw.cpp:116-127,155-170), found by a group of 8 lanes per pair reading the
 * captured last row together; a pair whose whole path is one diagonal (sw_trace.h clean_path) gets its result written
 * here, the others are appended to the worklist launch_traceback walks.  Must precede launch_traceback on the stream.
 */
cudaError_t launch_start(const TraceArgs &a, cudaStream_t stream);
/* has_short / has_long: the chunk has units of lane groups of 8 / whole-warp units; has_wide: some pair has MGLSW_HOP_WIDE_MIN_STRIPS
   column strips or more (a pipelined fill launch): the kernel with the four-block hop.  MGLSW_TRACE_ROUNDS + 1 launches for the
   short pairs, one for the long ones */
cudaError_t launch_traceback(const TraceArgs &a, bool has_short, bool has_long, bool has_wide, cudaStream_t stream);
/* opt-in: SOFTCLIP / IGNORE pairs whose alt occurs verbatim in the ref take GATK's wrapper answer (last occurrence, all M);
   runs after launch_traceback on the same stream and overwrites offsets / n_ops / scratch of those pairs */
cudaError_t launch_exact_match(const TraceArgs &a, cudaStream_t stream);
/* score-only runs: the traceback's start cell (sw.cpp:155-170) and its H, nothing else, by the same lane groups (needs units, strategy, c16, c32, colbest, lastrow, score3) */
cudaError_t launch_endpoints(const TraceArgs &a, cudaStream_t stream);

/* two-level exclusive scan of n_ops: starts[p] = offset inside p's tile of 1024 pairs, tile_sum[t] = offset of tile t,
   starts[n] = total.  tile_sum needs ceil(n / 1024) words. */
cudaError_t launch_scan_counts(const u32 *n_ops, u32 *starts, u32 *tile_sum, u32 n, cudaStream_t stream);

/* ops_out[tile_off[p / 1024] + starts[p] ..] = scratch[ops_end[p] - n_ops[p] .. ops_end[p]) */
cudaError_t launch_compact(const u32 *scratch, const u64 *ops_end, const u32 *n_ops, const u32 *starts, const u32 *tile_off,
                           u32 *ops_out, u32 n, cudaStream_t stream);

/* text CIGARs on the device: characters per pair, then (after launch_scan_counts over them) the characters themselves */
cudaError_t launch_text_len(const u32 *scratch, const u64 *ops_end, const u32 *n_ops, u32 *n_chars, u32 n, cudaStream_t stream);
cudaError_t launch_text_format(const u32 *scratch, const u64 *ops_end, const u32 *n_ops, const u32 *starts, const u32 *tile_off,
                               char *text_out, u32 n, cudaStream_t stream);

/* integer-pipe microbenchmarks (int_peak.cu): ops per second of one instruction mix */
cudaError_t run_int_peak(int which, int device, double *ops_per_s, double *sm_mhz);
int int_peak_count();
const char *int_peak_name(int which);

}  // namespace mglsw
#endif

/*
 * sw_plan.h -- host batcher, planning half (pure C++, no CUDA calls).
 *
 * Takes a chunk of pairs (the caller's arena offsets and lengths) and decides
 * how the GPU will process them: which arithmetic each pair can use, which
 * pairs share a register as 16-bit halves, how pairs are bucketed by length
 * into lane groups / warps / launches, and where each warp's trace, carry and
 * last-row words live.  The reference has nothing like this (its JNI is one
 * pair per call, ...SmithWaterman.cpp:44-71); it plays the role of align_avx's
 * per-call setup (sw_avx.cpp:6-60) for a whole batch.
 */
#ifndef MGL_B200_SW_PLAN_H
#define MGL_B200_SW_PLAN_H
#include <vector>
#include "sw_common.h"
#include "sw_config.h"

namespace mglsw {

struct PairIn {
    u64 ref_off, alt_off;
    u32 tl, ql;
    int strategy;
};

struct Launch {
    int kind;            /* KIND_S16X2 / KIND_S32 */
    int G, K;
    int pipe;            /* G == 32 only: one unit per block, its strips pipelined over the block's warps */
    u32 nstrips;
    u32 first_unit;      /* index into ChunkPlan::units */
    u32 n_units;         /* multiple of 32 / G (tail padded with inactive units) */
    u32 n_warps;
    u64 cells;           /* real DP cells in this launch */
};

struct ChunkPlan {
    std::vector<UnitDesc> units;
    std::vector<Launch> launches;
    std::vector<u32> pair_unit;     /* per chunk-local pair: unit index * 2 + half */
    std::vector<u64> ops_end;       /* per pair: one-past-the-end word of its CIGAR scratch */
    u32 n_units = 0;                /* unit descriptors of the chunk, padding included (== units.size() when planned on the host; */
                                    /* a chunk planned on the device has its descriptors in HBM only and `units` stays empty)     */
    u64 ckpt_words = 0;             /* column + row checkpoints of every warp */
    u64 carry_entries = 0;
    u64 lastrow_words = 0;
    u64 scratch_words = 0;
    u64 cells = 0;                  /* sum tl * ql */
    u64 pairs_s16 = 0, pairs_s32 = 0;
    int s16_bias = 0;
    /* scratch of plan_chunk, kept between calls so that a planner thread does not reallocate per chunk */
    std::vector<u64> full_key, sort_key, sort_tmp;
    std::vector<u32> sort_idx, sort_idx_tmp, sort_hist;
    std::vector<unsigned char> cand;
    void reset()                     /* empty, capacity kept */
    {
        units.clear(); launches.clear(); pair_unit.clear(); ops_end.clear(); n_units = 0;
        ckpt_words = carry_entries = lastrow_words = scratch_words = cells = pairs_s16 = pairs_s32 = 0;
        s16_bias = 0;
    }
};

/* limits a score set must respect for either arithmetic (checked once per batch) */
enum PlanStatus { PLAN_OK = 0, PLAN_BAD_SCORES = 1, PLAN_BAD_PAIR = 2 };

PlanStatus check_scores(const Scores &s);

/*
 * Can this pair be filled with unsigned 16-bit halves?  lo/hi receive
 * conservative bounds of every value a cell can hold (true scores, unbiased).
 */
bool s16_bounds(const Scores &s, u32 tl, u32 ql, long long &lo, long long &hi);

/* Plan one chunk.  allow_s16 = false forces the 32-bit kernels (tests use it). */
PlanStatus plan_chunk(const PairIn *pairs, u32 n, const Scores &s, bool allow_s16, ChunkPlan &out);

}  // namespace mglsw
#endif

/*
 * sw_engine.h -- host batcher, execution half: device buffers, streams and the
 * chunk pipeline of one GPU.  (The planning half is sw_plan.h.)
 *
 * A batch is cut into chunks of consecutive pairs.  Each chunk is planned on
 * the host while the previous one runs, its descriptors and the arena bytes it
 * touches go up on the chunk's stream, then: fill launches (one per bucket) ->
 * traceback -> scan -> compact -> results down.  Two work slots (trace, carry,
 * scratch, result buffers) alternate so copies of one chunk overlap kernels of
 * the other.  There is no CPU fallback: every error is returned to the caller.
 */
#ifndef MGL_B200_SW_ENGINE_H
#define MGL_B200_SW_ENGINE_H
#include <cuda_runtime.h>

#include <cstdint>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "sw_fill.h"
#include "sw_plan.h"
#include "sw_plan_dev_api.h"

/* work slots: chunks in flight at once (each owns checkpoint, scratch and result buffers and two streams) */
#define MGLSW_SLOTS 4
/* side streams of a slot: a chunk's fill launches (one per (arithmetic, G, K) bucket) are independent, so they are
   spread over these and run side by side; a bucket of a few long pairs no longer holds the whole GPU */
#define MGLSW_SIDE 7

namespace mglsw {

struct BatchIn {
    int64_t n;
    const uint8_t *arena;
    uint64_t arena_bytes;
    const uint64_t *ref_off, *alt_off;
    const uint32_t *ref_len, *alt_len;
    const int32_t *strategies;   /* per pair, or NULL */
    int strategy;                /* used when strategies == NULL */
    Scores scores;
    bool text = false;           /* CIGARs formatted on the device: BatchOut::text instead of ::ops */
    bool score_only = false;     /* no traceback: BatchOut::score3 only */
};

struct BatchOut {
    int32_t *offsets;            /* n */
    uint64_t *cigar_start;       /* n + 1 */
    uint32_t *ops;               /* ops_cap packed elements */
    uint64_t ops_cap;
    uint64_t ops_total;          /* out */
    char *text = nullptr;        /* text runs: text_cap characters, pair p's CIGAR is text[cigar_start[p] .. cigar_start[p+1]), no NUL; */
    uint64_t text_cap = 0;       /* ops_total then counts characters                                                             */
    int32_t *score3 = nullptr;   /* score-only runs: n x {score, end_ref, end_alt}; the other arrays are not touched */
    bool keep_first_start = false; /* cigar_start[0] is not written (it is the last entry of the previous shard of a multi-GPU call) */
};

struct BatchStats {              /* filled by every run */
    uint64_t cells, pairs, pairs_s16, pairs_s32;
    uint64_t fill_launches, other_launches, chunks;
    uint64_t trace_bytes;        /* checkpoint words written * 4 */
    uint64_t h2d_bytes, d2h_bytes;
    double fill_ms, trace_ms, post_ms, device_ms;   /* device_ms etc. only from run_resident */
};

/* growable device / pinned buffers */
struct DevBuf {
    void *p = nullptr; size_t cap = 0;
    cudaError_t reserve(size_t bytes);
    void release();
};
struct PinBuf {
    void *p = nullptr; size_t cap = 0;
    cudaError_t reserve(size_t bytes);
    void release();
};

/* a planned chunk with its descriptors on the device */
struct ChunkDev {
    ChunkPlan plan;
    int64_t first = 0; u32 n = 0;
    DevBuf units, pair_unit, ops_end, strategy;
    PinBuf h_units, h_pair_unit, h_ops_end, h_strategy;
    u64 arena_lo = 0, arena_hi = 0;      /* byte range of the caller's arena the chunk's pairs touch */
    bool packed_ok = false;              /* h_packed holds [arena_lo & ~15, arena_hi) in the 2-bit transport form (sw_pack.h) */
    DevBuf packed;
    PinBuf h_packed;
    std::vector<PairIn> pairs_scratch;
    /* what the kernels of this chunk read, whichever planner produced it */
    const UnitDesc *d_units = nullptr;
    const u64 *d_ops_end = nullptr;
    const int *d_strategy = nullptr;
    /* device planner (sw_plan_dev.cu): the chunk's slice of the pair table, staged in pinned memory and uploaded;
       work buffer; the summary the host reads back; the event that says the summary is home */
    DevBuf table, work;
    PinBuf h_table, h_sum;
    PinBuf h_arena;                      /* a pageable caller arena: the chunk's byte range, staged here by a few threads */
    PlanDevLayout lay{};
    cudaEvent_t ev_plan = nullptr;
    cudaEvent_t ev_table = nullptr;      /* the chunk's pair table is on the device */
    cudaEvent_t ev_arena = nullptr;      /* the chunk's arena bytes are on the device (sent ahead by the staging thread) */
    bool arena_ahead = false;
    cudaEvent_t ev_dbg[3] = {};          /* MGL_SW_TIMELINE: timing events around the planner's upload and kernels */
    void release();
};

struct WorkSlot {
    cudaStream_t stream = nullptr;       /* uploads and fill launches                                          */
    cudaStream_t post = nullptr;         /* traceback, scan, compact, downloads: higher priority, so that the   */
                                         /* latency-bound traceback of one chunk slips in between the blocks of */
                                         /* the other slot's fill instead of queueing behind them               */
    cudaStream_t side[MGLSW_SIDE] = {};  /* further fill launches of the same chunk, forked from / joined into `stream` */
    cudaEvent_t ev_fork = nullptr, ev_side[MGLSW_SIDE] = {};
    cudaEvent_t ev_fill = nullptr, ev_post = nullptr, ev_counts = nullptr, ev_done = nullptr;
    DevBuf ckpt, carry, lastrow, colbest, scratch, offsets, n_ops, starts, tiles, ops, score3, n_chars;
    DevBuf startcell, worklist, wstate;  /* launch_start -> launch_traceback (TraceArgs): counters in the first MGLSW_WC_WORDS words of worklist */
    PinBuf h_n_ops, h_total;
    PinBuf h_res, h_ops;                 /* results on their way to PAGEABLE caller arrays: offsets / scores, and CIGAR elements / text */
    void release();
};

/*
 * A few helper threads that stay: run(T, fn) calls fn(0) on the caller and fn(1) .. fn(T - 1) on them and returns when
 * all are done.  Creating and joining fifteen std::threads costs 0.5 ms on these boxes -- as much as the work they were
 * created for (the pair-table scan of a 1 M-pair batch) -- so the threads are created once and sleep in between (a
 * condition variable: nothing spins after the call and takes a core from the planner, as an OpenMP team would).
 */
class HelperPool {
public:
    ~HelperPool();
    void run(int T, const std::function<void(int)> &fn);
private:
    void worker();
    std::vector<std::thread> th_;
    std::mutex mu_, run_mu_;
    std::condition_variable cv_, done_cv_;
    const std::function<void(int)> *task_ = nullptr;
    int next_ = 0, total_ = 0, pending_ = 0;
    bool stop_ = false;
};

class Engine {
public:
    explicit Engine(int device);
    ~Engine();
    int device() const { return device_; }
    const std::string &last_error() const { return err_; }

    /* host buffers in, host buffers out; returns 0 or an MGL_SW_E_* code */
    int align_batch(const BatchIn &in, BatchOut &out, BatchStats *stats);

    /* device-resident mode (bench `value`): upload + plan once, then time kernels only */
    int prepare_resident(const BatchIn &in);
    int run_resident(BatchStats *stats);
    /* results of the last run_resident, brought down for checking */
    int fetch_resident(BatchOut &out);
    void drop_resident();

    bool allow_s16 = true;
    bool overlap = true;         /* run_resident: alternate the two slots so traceback(c) overlaps fill(c + 1);
                                    false = one chunk at a time with per-kernel event timing */
    u32 chunk_pairs = 1u << 16;  /* measured: 64 Ki pairs per chunk with four slots is within 1 % of the best device rate and plans in < 5 ms */
    int plan_threads = 0;        /* host planner threads of align_batch; 0: min(8, hardware threads / 2) */
    int scan_threads = 0;        /* short-lived threads of align_batch's pair-table scan; 0: min(8, hardware threads / 2) */
    bool exact_shortcut = false; /* SOFTCLIP / IGNORE pairs whose alt occurs verbatim in the ref answer as GATK's SWNativeAlignerWrapper does (sw_trace_kernel.cu) */
    int copy_threads = 8;        /* threads that stage a PAGEABLE caller arena into pinned memory, chunk by chunk (0: leave it to the driver) */
    int slots_opt = 0;           /* work slots per batch: 0 = two for batches of long pairs, four otherwise; 2..4 forces */
    int host_plan = 0;           /* 1: chunks are planned by host threads (sw_plan.cpp), 0: by device kernels (sw_plan_dev.cu) */
    int pipe_wide = 1;           /* pipelined long-pair launches on eight warps per pair: 0 never, 1 when the launch has few pairs, 2 always */
    u64 trace_budget_bytes = 0;  /* 0: derived from free memory */
    int pack_arena = 0;          /* align_batch: all-ACGT byte ranges cross the host link at 2 bits per base (sw_pack.h): */
                                 /* 0 never, 1 chunks whose planner is pack_lead chunks ahead of the uploader, 2 every chunk. */
                                 /* Off by default: on this pool's boxes the host cores and memory, not the link, are the   */
                                 /* shared resource (profiles/r03c, r03e: 8 GPUs 8 640 GCUPS raw, 7 875 / 6 255 with 1 / 2)  */
    int pack_lead = MGLSW_PACK_LEAD;
    int pack_threads = 0;        /* threads the device planner's staging thread packs a chunk with (0: two) */

private:
    int fail(int code, const std::string &msg);
    int cuda_fail(cudaError_t e, const char *what);
    int plan_next(const BatchIn &in, int64_t first, ChunkDev &cd);
    /* device planner: stage + upload the pair-table slice and queue the planning kernels on plan_stream_;
       then (later) wait for the summary and turn it into cd.plan */
    int issue_plan_dev(const BatchIn &in, int64_t first, int64_t take, ChunkDev &cd, BatchStats &bs, std::string &err, bool upload_arena, bool want_pack);
    int collect_plan_dev(ChunkDev &cd);
    int64_t chunk_take(const BatchIn &in, int64_t first, u64 limit, u32 max_pairs, const u32 *est256) const;
    int plan_range(const BatchIn &in, int64_t first, int64_t take, ChunkDev &cd, std::string &err) const;
    int pack_chunk(const BatchIn &in, ChunkDev &cd, std::string &err) const;
    int pack_chunk_mt(const BatchIn &in, ChunkDev &cd, int T, std::string &err);
    int upload_chunk(const BatchIn &in, ChunkDev &cd, cudaStream_t st, bool upload_arena, BatchStats &bs);
    int upload_arena_range(const BatchIn &in, ChunkDev &cd, cudaStream_t st, bool upload_arena, BatchStats &bs);
    int run_chunk(ChunkDev &cd, WorkSlot &ws, BatchStats &bs, bool timed, bool score_only = false, bool text = false);
    int ensure_slot(WorkSlot &ws, const ChunkPlan &plan, u32 n, bool score_only = false, bool text = false);
    u64 budget();
    void choose_slots(u64 total_est, int64_t n);
    int nslots_ = MGLSW_SLOTS;   /* work slots of the current batch */
    u32 chunk_cap(u64 total_est, int64_t n) const;
    u32 chunk_cap_ = 0;          /* prepare_resident: pairs per chunk of the resident batch */

    int device_;
    u64 budget_auto_ = 0;        /* budget() when trace_budget_bytes == 0, fixed at first use */
    std::string err_;
    DevBuf arena_;               /* device copy of bytes [arena_base_, ..) of the caller's arena, padded on both sides (sw_fill.h) */
    u64 arena_base_ = 0;
    HelperPool pool_;               /* the scan of a batch's pair table; copies and packing of a chunk's arena bytes */
    bool arena_pageable_ = false;   /* the current batch's arena is ordinary (unregistered) host memory */
    bool res_pageable_ = false, ops_pageable_ = false;   /* ... its offsets (or scores) / its CIGAR array are */
    /* where caller offset 0 would sit: device code adds the caller's own offsets (only [arena_base_, ..) is backed) */
    const uint8_t *arena_dev() const { return (const uint8_t *)((uintptr_t)arena_.p + MGLSW_ARENA_FRONT_PAD - (uintptr_t)arena_base_); }
    int reserve_arena(const BatchIn &in);
    int reserve_arena(u64 base, u64 span, u32 max_tl);
    std::vector<u64> estpre_;    /* align_batch: prefix sums of the per-pair checkpoint estimate */
    WorkSlot slots_[MGLSW_SLOTS];
    cudaStream_t plan_stream_ = nullptr;     /* device planner: high priority, runs ahead of the fills */
    cudaStream_t upload_stream_ = nullptr;   /* device planner mode: arena bytes of the chunks ahead */
    unsigned wait_flags_ = 0;                /* flags of the events the host waits on (spinning or sleeping) */
    std::vector<ChunkDev *> ring_;           /* align_batch: staging entries the planner threads fill */
    std::vector<ChunkDev *> resident_;
    std::vector<DevBuf> resident_results_;   /* per chunk: offsets | starts | ops kept on device */
    Scores resident_scores_{};
    cudaEvent_t tev_[5 + MGLSW_SLOTS] = {};
};

}  // namespace mglsw
#endif

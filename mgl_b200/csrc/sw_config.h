/*
 * sw_config.h -- which fill kernels exist.  One kernel per (arithmetic, lane
 * group width G, registers per lane K); the planner only emits launches whose
 * (kind, G, K) is in this table, and sw_fill_dispatch.cu fails loudly otherwise.
 */
#ifndef MGL_B200_SW_CONFIG_H
#define MGL_B200_SW_CONFIG_H

#define MGLSW_KMAX 24            /* registers (alt columns) per lane, upper bound */
#define MGLSW_G_SHORT 8          /* lane group for alt length <= 8 * KMAX         */
#define MGLSW_G_LONG 32          /* whole warp per pair, strip-mined              */
#define MGLSW_FILL_THREADS 128   /* 4 warps per block                             */
#define MGLSW_PIPE_WIDE_WARPS 8    /* warps per pair of the latency-shaped pipelined kernel (launches of few long pairs) */
#define MGLSW_PIPE_WIDE_MAX_UNITS 160  /* a chunk with at most this many pipelined pairs uses it */
#ifndef MGLSW_TRACE_ROUNDS
#define MGLSW_TRACE_ROUNDS 5        /* rebuilt blocks a short pair gets in the round kernels, one launch each, before a lane takes it to the end (sw_trace_kernel.cu) */
#endif
#define MGLSW_TRACE_ROUNDS_MIN_UNITS 1024   /* chunks of at most this many units skip the round kernels: one traceback launch */
#define MGLSW_DEV_PLAN_MIN_PAIRS 1024   /* batches of at most this many pairs are planned on the host (sw_engine.cu) */
#define MGLSW_WC_LONG 14           /* work_count word of the whole-warp units' list */
#define MGLSW_WC_WORDS 16
#define MGLSW_PIPE_MIN_STRIPS 4   /* pairs of at least this many strips run one per block, strips pipelined over its warps */
#define MGLSW_SPIN_MIN_CPUS_PER_GPU 0 /* below this many hardware threads per visible GPU the collecting thread sleeps on its events (0: always spin) */
#define MGLSW_PACK_LEAD 6          /* chunks a planner must be ahead of the uploader to spend time packing its chunk */
#define MGLSW_PACK_BACKLOG 2        /* device planning, pack_arena 1: uploads that must be waiting on the link before a chunk is packed */
#define MGLSW_PACK_FIRST_CHUNK 3    /* ... and never one of the first chunks of a batch */
#define MGLSW_PACK_MIN_BYTES 65536 /* a chunk's arena byte range below this is uploaded raw (one copy, no unpack launch) */

/* X(kind_type, G, K) for every instantiated kernel */
#define MGLSW_FOR_K_1_24(X, T, G) \
    X(T, G, 1) X(T, G, 2) X(T, G, 3) X(T, G, 4) X(T, G, 5) X(T, G, 6) X(T, G, 7) X(T, G, 8) \
    X(T, G, 9) X(T, G, 10) X(T, G, 11) X(T, G, 12) X(T, G, 13) X(T, G, 14) X(T, G, 15) X(T, G, 16) \
    X(T, G, 17) X(T, G, 18) X(T, G, 19) X(T, G, 20) X(T, G, 21) X(T, G, 22) X(T, G, 23) X(T, G, 24)
#define MGLSW_FOR_K_7_24(X, T, G) \
    X(T, G, 7) X(T, G, 8) \
    X(T, G, 9) X(T, G, 10) X(T, G, 11) X(T, G, 12) X(T, G, 13) X(T, G, 14) X(T, G, 15) X(T, G, 16) \
    X(T, G, 17) X(T, G, 18) X(T, G, 19) X(T, G, 20) X(T, G, 21) X(T, G, 22) X(T, G, 23) X(T, G, 24)

#endif

/*
 * mgl_sw.h -- C ABI of libmgl_sw.so, the B200 Smith-Waterman affine-gap aligner.
 *
 * Drop-in boundary.  The library exports the three JNI symbols the reference's
 * libmgl_sw.so exports (src/main/native/mgl_sw/
 * com_microsoft_mgl_smithwaterman_MicrosoftSmithWaterman.h:15,23,31), unchanged
 * in name, signature and behaviour, and adds underneath them the batched entry
 * points below, which are what reaches the GPU.  Nothing here takes or returns
 * a torch / CUDA type: plain pointers and sizes only.
 *
 * Semantics of one alignment are the reference's (sw.cpp:5-272): scores are
 * sign-normalised the way alignNative does (...SmithWaterman.cpp:52-55),
 * strategy is one of 1 SOFTCLIP, 2 INDEL, 4 LEADING_INDEL, 8 IGNORE
 * (sw_common.h:22-25), the result is the alignment offset and the CIGAR.
 * There is NO CPU fallback: without a usable B200 every call returns an error.
 */
#ifndef MGL_SW_H
#define MGL_SW_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MGL_SW_API __attribute__((visibility("default")))

/* status codes */
#define MGL_SW_OK               0
#define MGL_SW_E_INVALID       -1   /* NULL / negative / inconsistent argument                 */
#define MGL_SW_E_SCORES        -2   /* score set outside what 32-bit cells can represent       */
#define MGL_SW_E_PAIR          -3   /* empty or over-long sequence (the reference leaves empty
                                       inputs undefined, ...SmithWaterman.cpp:48-49)           */
#define MGL_SW_E_CUDA          -4   /* CUDA runtime error; see mgl_sw_last_error               */
#define MGL_SW_E_OPS_CAPACITY  -5   /* ops_out too small; *ops_total holds the needed count    */
#define MGL_SW_E_NO_DEVICE     -6   /* no CUDA device / extension unusable                     */

/* overhang strategies, sw_common.h:22-25 */
#define MGL_SW_SOFTCLIP       1
#define MGL_SW_INDEL          2
#define MGL_SW_LEADING_INDEL  4
#define MGL_SW_IGNORE         8

/* packed CIGAR element: (length << 4) | op, op in BAM numbering */
#define MGL_SW_OP_M 0
#define MGL_SW_OP_I 1
#define MGL_SW_OP_D 2
#define MGL_SW_OP_S 4

typedef struct mgl_sw_handle mgl_sw_handle;

/* One batch.  Sequences are raw bytes (compared for equality as bytes, sw.cpp:55) in one arena. */
typedef struct mgl_sw_batch {
    int64_t n;                    /* pairs                                                   */
    const uint8_t *arena;         /* sequence bytes (pinned memory makes the copy asynchronous) */
    uint64_t arena_bytes;
    const uint64_t *ref_off;      /* n: byte offset of each ref (target)                     */
    const uint32_t *ref_len;      /* n                                                       */
    const uint64_t *alt_off;      /* n: byte offset of each alt (query)                      */
    const uint32_t *alt_len;      /* n                                                       */
    const int32_t *strategies;    /* n, or NULL: `strategy` applies to every pair            */
    int32_t strategy;
    int32_t match, mismatch, gap_open, gap_extend;   /* as GATK passes them, any sign         */
} mgl_sw_batch;

typedef struct mgl_sw_result {
    int32_t *offsets;             /* n: alignment offsets (alignNative's return value)        */
    uint64_t *cigar_start;        /* n + 1: pair p's elements are ops[cigar_start[p] .. cigar_start[p+1]) */
    uint32_t *ops;                /* ops_cap packed elements, left to right                   */
    uint64_t ops_cap;
    uint64_t ops_total;           /* out: elements produced (needed, on MGL_SW_E_OPS_CAPACITY) */
} mgl_sw_result;

typedef struct mgl_sw_stats {
    uint64_t cells, pairs, pairs_s16, pairs_s32;
    uint64_t fill_launches, other_launches, chunks;
    uint64_t trace_bytes, h2d_bytes, d2h_bytes;
    double fill_ms, trace_ms, post_ms, device_ms;
} mgl_sw_stats;

/* lifecycle: a handle drives one GPU, or a list of them; calls on one handle are serialised internally */
MGL_SW_API int mgl_sw_create(int device, mgl_sw_handle **out);
/*
 * One handle over several GPUs of the box (SURVEY.md 8b "handle create/destroy (device list, streams)", 8e): every
 * batch call on it cuts the batch into n_devices contiguous ranges of near-equal DP cells, runs range k on
 * devices[k] (one host thread and one engine per device, nothing exchanged between GPUs: pairs are independent)
 * and gathers the results in input order.  The same device may be listed more than once.  The resident_* entries
 * use devices[0] only.
 */
MGL_SW_API int mgl_sw_create_multi(const int *devices, int n_devices, mgl_sw_handle **out);
MGL_SW_API int mgl_sw_handle_devices(mgl_sw_handle *h, int *devices, int cap);   /* returns the device count */
MGL_SW_API void mgl_sw_destroy(mgl_sw_handle *h);
/*
 * The process-wide handle behind the entry points that carry no handle (mgl_sw_align_one and the JNI symbols):
 * created by the first caller over every visible GPU, or the comma-separated MGL_SW_DEVICES list, or the one device
 * MGL_SW_DEVICE names.  mgl_sw_default_handle pins it (a reference count), mgl_sw_default_release unpins it;
 * mgl_sw_shutdown -- what doneNative calls, MicrosoftSmithWaterman.java:89-92 -- waits for pinned users, then
 * destroys it and frees its device memory; a later call creates it afresh.
 */
MGL_SW_API int mgl_sw_default_handle(mgl_sw_handle **out);
MGL_SW_API void mgl_sw_default_release(void);
MGL_SW_API void mgl_sw_shutdown(void);
MGL_SW_API const char *mgl_sw_last_error(mgl_sw_handle *h);
MGL_SW_API int mgl_sw_device_count(void);
MGL_SW_API const char *mgl_sw_version(void);

/* tuning knobs: "allow_s16" 0/1, "chunk_pairs", "trace_budget_mb", "plan_threads", "pack_arena" 0/1/2 (all-ACGT byte
   ranges of the arena cross the host link at 2 bits per base and are expanded on the device: never / when it hides behind the
   link (host planning: chunks whose planner thread is "pack_lead" chunks ahead of the upload; device planning: while uploads are queued
   and chunks' bytes are late) / always; default 0: it pays only where the link, not the
   host's cores and memory, is the slower side), "pipe_wide" 0/1/2 (long-pair
   kernel on eight warps per pair: never / for launches of few pairs / always), "copy_threads" (threads that stage a PAGEABLE arena into pinned
   memory chunk by chunk; default 8 or half the hardware threads, 0 leaves it to the driver), "exact_match_shortcut" 0/1
   (default 0: the reference's own output; 1: a SOFTCLIP / IGNORE pair whose alt occurs verbatim in its ref returns
   what GATK's SWNativeAlignerWrapper returns without calling the binding -- offset = last occurrence, CIGAR all M --
   which the reference leaves to that wrapper, ...MicrosoftSmithWaterman.cpp:60), "slots" 0/2/3/4 (work slots, i.e. chunks
   in flight: 0 = two for batches of long pairs, four otherwise), "host_plan" 0/1 (chunks planned by
   device kernels / by host threads), "pack_threads" (threads the 2-bit transport packs a chunk with under device planning; 0 = two), "overlap" 0/1 (resident_run: 1 = chunks
   alternate between two slots, whole-batch time only; 0 = one chunk at a time with per-kernel times) */
MGL_SW_API int mgl_sw_set_option(mgl_sw_handle *h, const char *name, int64_t value);

/* the batched aligner: host buffers in, host buffers out */
MGL_SW_API int mgl_sw_align_batch(mgl_sw_handle *h, const mgl_sw_batch *batch, mgl_sw_result *result, mgl_sw_stats *stats);

/*
 * The same batch with the CIGARs formatted on the device: pair p's text is
 * text[text_start[p] .. text_start[p+1]), exactly the characters alignNative writes to cigarBuffer
 * (to_string(length) + state per element, sw.cpp:251-252; no NUL, ...SmithWaterman.cpp:65).
 * MGL_SW_E_OPS_CAPACITY when text_cap is too small: text_total holds the characters needed.
 */
typedef struct mgl_sw_text_result {
    int32_t *offsets;             /* n                                                        */
    uint64_t *text_start;         /* n + 1                                                    */
    char *text;                   /* text_cap characters                                      */
    uint64_t text_cap;
    uint64_t text_total;          /* out */
} mgl_sw_text_result;
MGL_SW_API int mgl_sw_align_batch_text(mgl_sw_handle *h, const mgl_sw_batch *batch, mgl_sw_text_result *result, mgl_sw_stats *stats);

/*
 * Score only (no caller in the reference today): the fill without checkpoints and no
 * traceback.  score3[3p .. 3p+2] = { H of the cell the traceback would start from, its row (ref
 * position, 1-based), its column (alt position) }: the corner for INDEL, the last-column maximum for
 * LEADING_INDEL, the better of last column and last row otherwise (ScoreMax, sw_common.h:36-40, as used by
 * sw.cpp:155-170).  Same inputs, errors and threading as mgl_sw_align_batch.
 */
MGL_SW_API int mgl_sw_score_batch(mgl_sw_handle *h, const mgl_sw_batch *batch, int32_t *score3, mgl_sw_stats *stats);

/* device-resident variant for kernel timing: inputs uploaded once, kernels timed with CUDA events */
MGL_SW_API int mgl_sw_resident_prepare(mgl_sw_handle *h, const mgl_sw_batch *batch);
MGL_SW_API int mgl_sw_resident_run(mgl_sw_handle *h, mgl_sw_stats *stats);
MGL_SW_API int mgl_sw_resident_fetch(mgl_sw_handle *h, mgl_sw_result *result);
MGL_SW_API void mgl_sw_resident_drop(mgl_sw_handle *h);

/* one pair, text CIGAR, exactly what alignNative produces (batch of one on a process-wide handle) */
MGL_SW_API int mgl_sw_align_one(const uint8_t *ref, int32_t ref_len, const uint8_t *alt, int32_t alt_len,
                                int32_t match, int32_t mismatch, int32_t gap_open, int32_t gap_extend,
                                int32_t strategy, char *cigar_out, int32_t cigar_cap, int32_t *cigar_len,
                                int32_t *offset);

/* packed elements -> "3M2D5M" (sw.cpp:251-252); returns characters written or -1 */
MGL_SW_API int mgl_sw_format_cigar(const uint32_t *ops, int64_t n_ops, char *out, int64_t out_cap);

/* pinned host memory for arenas / results */
MGL_SW_API void *mgl_sw_host_alloc(uint64_t bytes);
MGL_SW_API void mgl_sw_host_free(void *p);

/* host half of the 2-bit transport form (no GPU needed; tests): packs src[0, n) into dst, four bases per byte, base k in
   bits 2 (k % 4) of byte k / 4, code = (byte >> 1) & 3; dst holds ((n + 15) / 16) * 4 bytes.  Returns 1 when packed,
   0 when a byte outside ACGT makes the range travel raw, -1 on a NULL pointer or a dst_cap too small.
   `scalar` != 0 forces the table loop instead of AVX2 */
MGL_SW_API int mgl_sw_pack2bit_host(const uint8_t *src, uint64_t n, uint8_t *dst, uint64_t dst_cap, int scalar);

/* planning only (no GPU needed): how a batch would be bucketed; used by tests and tools */
MGL_SW_API int mgl_sw_plan_summary(const mgl_sw_batch *batch, int allow_s16, uint64_t *out8);

/* integer issue-rate microbenchmarks for the roofline denominator */
MGL_SW_API int mgl_sw_int_peak_count(void);
MGL_SW_API const char *mgl_sw_int_peak_name(int which);
MGL_SW_API int mgl_sw_int_peak(int device, int which, double *ops_per_s, double *sm_mhz);

#ifdef __cplusplus
}
#endif
#endif

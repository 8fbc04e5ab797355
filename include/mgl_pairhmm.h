/*
 * mgl_pairhmm.h -- C-ABI of the B200 PairHMM (libmgl_pairhmm.so), the library's second kernel
 * (SURVEY.md 8f row 3).  Replaces, behind the same JNI symbols, the reference's
 *   Java_com_microsoft_mgl_pairhmm_MicrosoftPairHmm_initNative / computeLikelihoodsNative / doneNative
 *   (src/main/native/mgl_pairhmm/com_microsoft_mgl_pairhmm_MicrosoftPairHmm.h:15-31, .cc:52-235)
 * and underneath them the forward algorithm of compute_prob_scalar.cc / compute_prob_avx{f,d}.cc.
 *
 * The job is the reference's: every read against every haplotype, log10 likelihood per pair, computed in
 * float first and again in double where the float result is below MIN_ACCEPTED = 1e-28f
 * (pairhmm_common.h:31, .cc:176-203), or in double throughout when use_double is set (.cc:165-171).
 * Plain pointers and sizes; host buffers in, host buffers out; no CPU fallback.
 */
#ifndef MGL_PAIRHMM_H
#define MGL_PAIRHMM_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define MGL_PH_API __attribute__((visibility("default")))
#else
#define MGL_PH_API
#endif

#define MGL_PH_OK            0
#define MGL_PH_E_INVALID    -1   /* NULL pointer, negative count, empty read or haplotype            */
#define MGL_PH_E_NO_DEVICE  -2   /* no CUDA device / not an sm_100 part / library built without it    */
#define MGL_PH_E_CUDA       -3   /* a CUDA call failed: mgl_pairhmm_last_error has the text           */
#define MGL_PH_E_TOO_LONG   -4   /* a haplotype too long for the shared-memory carry row (see below)  */

typedef struct mgl_pairhmm_handle mgl_pairhmm_handle;

typedef struct mgl_pairhmm_stats {
    uint64_t pairs, cells;            /* read x haplotype pairs; sum of readLen * hapLen                */
    uint64_t pairs_double;            /* pairs computed in double (escalated, or all with use_double)   */
    uint64_t launches;                /* kernels launched                                               */
    uint64_t h2d_bytes, d2h_bytes;
    double device_ms;                 /* CUDA events around the kernels of this call                    */
} mgl_pairhmm_stats;

MGL_PH_API int mgl_pairhmm_create(int device, mgl_pairhmm_handle **out);
MGL_PH_API void mgl_pairhmm_destroy(mgl_pairhmm_handle *h);
MGL_PH_API const char *mgl_pairhmm_last_error(mgl_pairhmm_handle *h);

/*
 * Buffers exactly as MicrosoftPairHmm.java hands them to computeLikelihoodsNative (.cc:84-122):
 *   lengths : int32 { readCount, readLen[readCount], hapCount, hapLen[hapCount] }
 *   reads   : per read  bases | base quals | insertion GOP | deletion GOP | GCP   (readLen bytes each, phred)
 *   haps    : haplotype bases back to back
 *   out     : double[readCount * hapCount], read-major: log10 likelihood
 * Reads of up to 256 bases (one band of 32 lanes x 8 rows) run against haplotypes of any length.  A call holding a
 * longer read keeps a carry row per haplotype column in shared memory and is then limited to haplotypes of 2 344
 * bases (220 KiB / 4 warps / 24 B): beyond that the call fails with MGL_PH_E_TOO_LONG -- through JNI a
 * RuntimeException, never a silent result.  (The reference has no such limit; GATK's reads and haplotypes sit far
 * inside it.)
 */
MGL_PH_API int mgl_pairhmm_compute(mgl_pairhmm_handle *h, const int32_t *lengths, const uint8_t *reads, const uint8_t *haps,
                                   int use_double, double *out, mgl_pairhmm_stats *stats);

/* the same job with the three inputs already on the device (bench `value`): upload once, then run kernels only */
MGL_PH_API int mgl_pairhmm_resident_prepare(mgl_pairhmm_handle *h, const int32_t *lengths, const uint8_t *reads, const uint8_t *haps);
MGL_PH_API int mgl_pairhmm_resident_run(mgl_pairhmm_handle *h, int use_double, mgl_pairhmm_stats *stats);
MGL_PH_API int mgl_pairhmm_resident_fetch(mgl_pairhmm_handle *h, double *out);

/*
 * The same job for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  NO COUNTERPART IN THE REFERENCE.
 *
 * Needs no handle and allocates nothing, neither device nor pinned memory: no cudaMalloc / cudaFree, no stream or
 * event creation.  Everything mgl_pairhmm_compute keeps in cudaMalloc'ed buffers -- the float pass's redo list and
 * its counter, the quality tables, the offset tables, the padded copies of reads and haplotypes, the likelihoods
 * before they are copied out -- lives in ONE file-scope static __device__ array of the library
 * (mgl_pairhmm_noalloc_scratch_bytes() bytes, about 3 MiB, held by every GPU that loads the library until exit).
 * Hence the limits, beyond which the call fails with MGL_PH_E_TOO_LARGE and touches nothing:
 *   reads <= MGL_PH_NOALLOC_MAX_READS, haplotypes <= MGL_PH_NOALLOC_MAX_HAPS, pairs <= MGL_PH_NOALLOC_MAX_PAIRS,
 *   read bases in total <= MGL_PH_NOALLOC_MAX_READ_BASES, haplotype bases in total <= MGL_PH_NOALLOC_MAX_HAP_BASES.
 *
 * lengths is host memory (as above); reads, haps and out are the caller's buffers in the layout above and may be
 * device pointers or host pointers (copied with cudaMemcpyDefault; pinned host memory keeps the copies asynchronous).
 * stream is a cudaStream_t passed as a pointer, NULL for the default stream: copies and kernels are enqueued on it
 * and, unless flags has MGL_PH_NOALLOC_NO_SYNC, waited for before the call returns.  The first call on a device
 * does a one-time synchronous set-up (symbol address, table upload): make it before a capture begins.
 * Every call shares the one array: host-side the calls are serialised, but two calls whose work OVERLAPS ON THE GPU
 * (different streams, or NO_SYNC calls on streams that run concurrently) race on it -- order them on one stream.
 * All reads run one kernel shape chosen from the longest read (mgl_pairhmm_compute picks a shape per read): results
 * agree with mgl_pairhmm_compute within the library's 1e-5 tolerance on the log10 likelihood.
 */
#define MGL_PH_E_TOO_LARGE  -5   /* job beyond the static scratch of mgl_pairhmm_compute_noalloc       */
#define MGL_PH_NOALLOC_MAX_READS       4096
#define MGL_PH_NOALLOC_MAX_HAPS        1024
#define MGL_PH_NOALLOC_MAX_PAIRS       65536
#define MGL_PH_NOALLOC_MAX_READ_BASES  262144
#define MGL_PH_NOALLOC_MAX_HAP_BASES   262144
#define MGL_PH_NOALLOC_NO_SYNC         1
MGL_PH_API int mgl_pairhmm_compute_noalloc(int device, const int32_t *lengths, const uint8_t *reads, const uint8_t *haps,
                                           int use_double, double *out, void *stream, int flags);
MGL_PH_API uint64_t mgl_pairhmm_noalloc_scratch_bytes(void);
MGL_PH_API const char *mgl_pairhmm_noalloc_last_error(void);   /* of the calling thread's last failed call */

#ifdef __cplusplus
}
#endif
#endif

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

#ifdef __cplusplus
}
#endif
#endif

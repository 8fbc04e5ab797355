/*
 * oracle/ref/ref_driver.cpp -- TEST INFRASTRUCTURE.
 * C entry points around the reference's own, unmodified C++ aligners
 * (align_scalar sw.cpp:258, align_avx sw_avx.cpp:6, and the JNI entry
 * ...MicrosoftSmithWaterman.cpp:44) so that tests and the CPU-baseline leg of
 * bench.py can execute the real reference.  Linked into oracle/_ref/libmgl_sw_ref.so
 * together with the reference sources compiled where they lie.
 */
#include <chrono>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "com_microsoft_mgl_smithwaterman_MicrosoftSmithWaterman.h"
#include "sw_scalar.h"
#include "sw_avx.h"

namespace {

void *identity_buffer_address(JNIEnv *, jobject buf) { return (void *)buf; }

struct FakeJvm {
    JNINativeInterface_ table;
    JNIEnv_ env;
    FakeJvm() { std::memset(&table, 0, sizeof table); table.GetDirectBufferAddress = identity_buffer_address; env.functions = &table; }
};

swParameters normalise(int match, int mismatch, int go, int ge)
{   /* same normalisation as ...MicrosoftSmithWaterman.cpp:52-55 */
    swParameters p;
    p.sc_match = match > 0 ? match : -match;
    p.sc_mismatch = mismatch < 0 ? mismatch : -mismatch;
    p.g_open = go > 0 ? go : -go;
    p.g_ext = ge > 0 ? ge : -ge;
    return p;
}

int run_one(int kind, const char *ref, int tl, const char *alt, int ql, int match, int mismatch,
            int go, int ge, int strategy, std::string *cigar)
{
    if (kind == 0) return align_scalar(ref, tl, alt, ql, normalise(match, mismatch, go, ge), strategy, cigar);
    if (kind == 1) return align_avx(ref, tl, alt, ql, normalise(match, mismatch, go, ge), strategy, cigar);
    /* kind 2: the reference's JNI dispatch rule (avx2 iff queryLength >= 8), without a JVM */
    if (__builtin_cpu_supports("avx2") && ql >= 8)
        return align_avx(ref, tl, alt, ql, normalise(match, mismatch, go, ge), strategy, cigar);
    return align_scalar(ref, tl, alt, ql, normalise(match, mismatch, go, ge), strategy, cigar);
}

}  // namespace

extern "C" {

/* kind: 0 align_scalar, 1 align_avx, 2 JNI dispatch rule.  Returns cigar length or -1 if cap too small. */
int mgl_ref_align(int kind, const uint8_t *ref, int tl, const uint8_t *alt, int ql,
                  int match, int mismatch, int go, int ge, int strategy,
                  char *cigar_out, int cigar_cap, int32_t *offset)
{
    std::string cigar;
    *offset = run_one(kind, (const char *)ref, tl, (const char *)alt, ql, match, mismatch, go, ge, strategy, &cigar);
    if ((int)cigar.size() >= cigar_cap) return -1;
    std::memcpy(cigar_out, cigar.data(), cigar.size());
    cigar_out[cigar.size()] = 0;
    return (int)cigar.size();
}

/* Through the reference's real exported JNI symbol with a table-backed JNIEnv (buffers = raw addresses). */
int mgl_ref_align_jni(const uint8_t *reads, int tl, int ql, int match, int mismatch, int go, int ge,
                      int strategy, char *cigar_buf)
{
    static FakeJvm jvm;
    return Java_com_microsoft_mgl_smithwaterman_MicrosoftSmithWaterman_alignNative(
        &jvm.env, nullptr, (jobject)reads, (jobject)cigar_buf, tl, ql, match, mismatch, go, ge, strategy);
}

/*
 * The reference's ScoreMax (sw_common.h:36-40) for one pair: its own calculateMatrix (sw.cpp:5-146) run the way
 * align_scalar does (sw.cpp:258-272), the traceback left out.  out6 = {mqe, mqe_t, max, max_t, max_q, seg_length}.
 * Pins the start cell and score of the score-only entry point.  Returns 0, -1 out of memory.
 */
int mgl_ref_scoremax(const uint8_t *ref, int tl, const uint8_t *alt, int ql,
                     int match, int mismatch, int go, int ge, int strategy, int32_t *out6)
{
    ScoreMax ez;
    const size_t cells = ((size_t)tl + 1) * ((size_t)ql + 1);
    int *bck = (int *)calloc(cells, sizeof(int));
    if (!bck) return -1;
    calculateMatrix((const char *)ref, tl, (const char *)alt, ql, bck, normalise(match, mismatch, go, ge), strategy, &ez);
    free(bck);
    out6[0] = ez.mqe; out6[1] = ez.mqe_t; out6[2] = ez.max; out6[3] = ez.max_t; out6[4] = ez.max_q; out6[5] = ez.seg_length;
    return 0;
}

int mgl_ref_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/*
 * Batch runner for the timed CPU baseline: one pair per task over `threads`
 * OpenMP threads.  pairs[k] = {ref_off, alt_off, ref_len, alt_len} into arena.
 * strategies may be NULL (then `strategy` applies to all).  Outputs (optional):
 * offsets[n], cigar text packed at cigars + k*cigar_stride (NUL terminated).
 * Returns wall seconds.
 */
double mgl_ref_align_batch(int kind, int64_t n, const uint8_t *arena, const uint64_t *ref_off,
                           const uint64_t *alt_off, const uint32_t *ref_len, const uint32_t *alt_len,
                           int match, int mismatch, int go, int ge, int strategy, const uint8_t *strategies,
                           int threads, int32_t *offsets, char *cigars, int cigar_stride)
{
    if (threads < 1) threads = 1;
    auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic, 16) num_threads(threads)
    for (int64_t k = 0; k < n; k++) {
        std::string cigar;
        int st = strategies ? strategies[k] : strategy;
        int off = run_one(kind, (const char *)arena + ref_off[k], (int)ref_len[k], (const char *)arena + alt_off[k],
                          (int)alt_len[k], match, mismatch, go, ge, st, &cigar);
        if (offsets) offsets[k] = off;
        if (cigars) {
            size_t len = cigar.size() < (size_t)cigar_stride - 1 ? cigar.size() : (size_t)cigar_stride - 1;
            std::memcpy(cigars + k * (int64_t)cigar_stride, cigar.data(), len);
            cigars[k * (int64_t)cigar_stride + len] = 0;
        }
    }
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

/*
 * Full-size parity without n x stride bytes of text: runs the reference on every pair (one pair per OpenMP task) and
 * compares with the caller's results in place -- offsets[k] and the packed elements (length << 4 | op, BAM numbering)
 * ops[cigar_start[k] .. cigar_start[k+1]), printed the way sw.cpp:251-252 prints them.  Returns the number of
 * differing pairs; the first `max_bad` differing indices land in bad_idx (ascending).  *secs = wall seconds.
 */
int64_t mgl_ref_check_batch(int kind, int64_t n, const uint8_t *arena, const uint64_t *ref_off,
                            const uint64_t *alt_off, const uint32_t *ref_len, const uint32_t *alt_len,
                            int match, int mismatch, int go, int ge, int strategy, const int32_t *strategies,
                            int threads, const int32_t *got_offsets, const uint64_t *cigar_start, const uint32_t *ops,
                            int64_t *bad_idx, int max_bad, double *secs)
{
    if (threads < 1) threads = 1;
    static const char kOp[] = "MIDNSHP=X";
    auto t0 = std::chrono::steady_clock::now();
    int64_t nbad = 0;
    std::vector<unsigned char> bad((size_t)n, 0);
#pragma omp parallel for schedule(dynamic, 16) num_threads(threads) reduction(+ : nbad)
    for (int64_t k = 0; k < n; k++) {
        std::string cigar, got;
        const int st = strategies ? strategies[k] : strategy;
        const int off = run_one(kind, (const char *)arena + ref_off[k], (int)ref_len[k], (const char *)arena + alt_off[k],
                                (int)alt_len[k], match, mismatch, go, ge, st, &cigar);
        for (uint64_t e = cigar_start[k]; e < cigar_start[k + 1]; e++) {
            const uint32_t op = ops[e] & 15u;
            got += std::to_string(ops[e] >> 4);
            got += op < 9 ? kOp[op] : '?';
        }
        if (off != got_offsets[k] || got != cigar) { bad[(size_t)k] = 1; nbad++; }
    }
    int w = 0;
    for (int64_t k = 0; k < n && w < max_bad; k++) if (bad[(size_t)k]) bad_idx[w++] = k;
    if (secs) *secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return nbad;
}

}  // extern "C"

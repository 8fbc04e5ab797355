/*
 * oracle/ref/pairhmm_driver.cpp -- TEST INFRASTRUCTURE.
 * C entry points around the reference's own, unmodified PairHMM (its JNI entry
 * com_microsoft_mgl_pairhmm_MicrosoftPairHmm.cc:77 with the AVX2 / scalar kernels behind it), so that
 * tests and the CPU-baseline leg of tools/bench_pairhmm.py can execute the real reference.  Linked into
 * oracle/_ref/libmgl_pairhmm_ref.so with the reference sources compiled where they lie; TBB is replaced by
 * the serial stand-in oracle/ref/tbb_stub (reads are independent), and this driver spreads the reads over
 * OpenMP threads instead, one JNI call per slice of reads.
 *
 * Buffer layout = what MicrosoftPairHmm.java hands to computeLikelihoodsNative (.cc:84-122):
 *   lengths : int32  { readCount, readLen[readCount], hapCount, hapLen[hapCount] }
 *   reads   : per read  bases | quals | insGOP | delGOP | GCP   (readLen bytes each)
 *   haps    : per haplotype its bases
 *   out     : double[readCount * hapCount], read-major
 */
#include <chrono>
#include <cstdint>
#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "com_microsoft_mgl_pairhmm_MicrosoftPairHmm.h"

namespace {
void *identity_buffer_address(JNIEnv *, jobject buf) { return (void *)buf; }
struct FakeJvm {
    JNINativeInterface_ table;
    JNIEnv_ env;
    FakeJvm() { std::memset(&table, 0, sizeof table); table.GetDirectBufferAddress = identity_buffer_address; env.functions = &table; }
};
}  // namespace

extern "C" {

/* one call of the reference's JNI entry on the calling thread */
void mgl_ref_pairhmm(const int32_t *lengths, const uint8_t *reads, const uint8_t *haps, int use_double, double *out)
{
    static FakeJvm jvm;
    Java_com_microsoft_mgl_pairhmm_MicrosoftPairHmm_initNative(&jvm.env, nullptr, (jboolean)(use_double != 0), 1);
    Java_com_microsoft_mgl_pairhmm_MicrosoftPairHmm_computeLikelihoodsNative(
        &jvm.env, nullptr, (jobject)lengths, (jobject)reads, (jobject)haps, (jobject)out);
}

/* the same job with the reads cut into slices, one JNI call per slice, over `threads` OpenMP threads; returns wall seconds */
double mgl_ref_pairhmm_threads(const int32_t *lengths, const uint8_t *reads, const uint8_t *haps, int use_double,
                               double *out, int threads, int reads_per_call)
{
    static FakeJvm jvm;
    Java_com_microsoft_mgl_pairhmm_MicrosoftPairHmm_initNative(&jvm.env, nullptr, (jboolean)(use_double != 0), 1);
    const int32_t R = lengths[0];
    const int32_t *rl = lengths + 1;
    const int32_t H = lengths[1 + R];
    const int32_t *hl = lengths + 2 + R;
    std::vector<int64_t> roff((size_t)R + 1, 0);
    for (int32_t r = 0; r < R; r++) roff[(size_t)r + 1] = roff[(size_t)r] + 5ll * rl[r];
    if (threads < 1) threads = 1;
    if (reads_per_call < 1) reads_per_call = 1;
    const int32_t calls = (R + reads_per_call - 1) / reads_per_call;
    auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
    for (int32_t c = 0; c < calls; c++) {
        const int32_t lo = c * reads_per_call, hi = lo + reads_per_call < R ? lo + reads_per_call : R;
        std::vector<int32_t> len;
        len.push_back(hi - lo);
        for (int32_t r = lo; r < hi; r++) len.push_back(rl[r]);
        len.push_back(H);
        for (int32_t h = 0; h < H; h++) len.push_back(hl[h]);
        Java_com_microsoft_mgl_pairhmm_MicrosoftPairHmm_computeLikelihoodsNative(
            &jvm.env, nullptr, (jobject)len.data(), (jobject)(reads + roff[(size_t)lo]), (jobject)haps, (jobject)(out + (int64_t)lo * H));
    }
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

int mgl_ref_pairhmm_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

}  // extern "C"

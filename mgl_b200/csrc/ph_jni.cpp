/*
 * ph_jni.cpp -- the three JNI symbols of the reference's libmgl_pairhmm.so with the reference's exact signatures
 * (src/main/native/mgl_pairhmm/com_microsoft_mgl_pairhmm_MicrosoftPairHmm.h:15,23,31), so the unmodified Java class
 * MicrosoftPairHmm binds to this library.
 *
 * Behaviour kept from ...MicrosoftPairHmm.cc:52-235: initNative(use_double, max_threads) records the precision mode
 * (max_threads has no meaning on a GPU); computeLikelihoodsNative takes the four direct buffers (IntBuffer lengths,
 * ByteBuffer reads, ByteBuffer haplotypes, DoubleBuffer likelihoods) and fills likelihoods[read * hapCount + hap].
 * The reference has no error channel.  MicrosoftPairHmm.java hands in a fresh zero-filled DoubleBuffer per call, so
 * returning quietly after a failure (no GPU, CUDA error, over-long haplotype, invalid input) would read as log10
 * likelihood 0.0 -- probability 1 -- for every pair.  A failure therefore fills the likelihoods with NaN and raises
 * java.lang.RuntimeException (jni_throw.h); it never computes on the CPU.  The GPU context is created by the first
 * compute call and released by doneNative.
 */
#include <jni.h>

#include <cmath>
#include <cstdio>
#include <mutex>
#include <string>

#include "../../include/mgl_pairhmm.h"
#include "jni_throw.h"

namespace {
bool g_use_double = false;
mgl_pairhmm_handle *g_handle = nullptr;
std::mutex g_mu;
}

extern "C" {

JNIEXPORT void JNICALL Java_com_microsoft_mgl_pairhmm_MicrosoftPairHmm_initNative(JNIEnv *, jclass, jboolean use_double, jint)
{
    g_use_double = use_double != 0;
}

JNIEXPORT void JNICALL Java_com_microsoft_mgl_pairhmm_MicrosoftPairHmm_computeLikelihoodsNative(
    JNIEnv *env, jobject, jobject lengthBuffer, jobject readsBuffer, jobject haplotypesBuffer, jobject likelihoodBuffer)
{
    int create_rc = MGL_PH_OK;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        if (!g_handle) {
            const int rc = mgl_pairhmm_create(0, &g_handle);
            if (rc != MGL_PH_OK) { g_handle = nullptr; create_rc = rc; }
        }
    }
    const int32_t *lengths = (const int32_t *)env->GetDirectBufferAddress(lengthBuffer);
    const uint8_t *reads = (const uint8_t *)env->GetDirectBufferAddress(readsBuffer);
    const uint8_t *haps = (const uint8_t *)env->GetDirectBufferAddress(haplotypesBuffer);
    double *out = (double *)env->GetDirectBufferAddress(likelihoodBuffer);
    /* whatever goes wrong below, the caller must not read its zero-filled buffer as results */
    auto poison = [&]() {
        if (!lengths || !out || lengths[0] < 0) return;
        const int64_t R = lengths[0];
        const int64_t H = lengths[1 + R];
        if (H < 0) return;
        for (int64_t k = 0; k < R * H; k++) out[k] = std::nan("");
    };
    if (create_rc != MGL_PH_OK) {
        poison();
        const std::string msg = "mgl_pairhmm computeLikelihoodsNative: no usable B200 (sm_100a) device (code " + std::to_string(create_rc) + ")";
        mgl_jni_throw(env, msg.c_str());
        return;
    }
    if (!lengths || !reads || !haps || !out) {
        mgl_jni_throw(env, "mgl_pairhmm computeLikelihoodsNative: every buffer must be direct");
        return;
    }
    mgl_pairhmm_handle *h;
    { std::lock_guard<std::mutex> lk(g_mu); h = g_handle; }
    const int rc = mgl_pairhmm_compute(h, lengths, reads, haps, g_use_double ? 1 : 0, out, nullptr);
    if (rc != MGL_PH_OK) {
        poison();
        const std::string msg = "mgl_pairhmm computeLikelihoodsNative: GPU computation failed (code " + std::to_string(rc) + ": " +
                                mgl_pairhmm_last_error(h) + "); likelihoods set to NaN";
        mgl_jni_throw(env, msg.c_str());
    }
}

/* MicrosoftPairHmm.done() -> doneNative: the GPU context and its buffers go; the next compute call creates them again */
JNIEXPORT void JNICALL Java_com_microsoft_mgl_pairhmm_MicrosoftPairHmm_doneNative(JNIEnv *, jobject)
{
    std::lock_guard<std::mutex> lk(g_mu);
    if (g_handle) { mgl_pairhmm_destroy(g_handle); g_handle = nullptr; }
}

}  // extern "C"

/*
 * ph_jni.cpp -- the three JNI symbols of the reference's libmgl_pairhmm.so with the reference's exact signatures
 * (src/main/native/mgl_pairhmm/com_microsoft_mgl_pairhmm_MicrosoftPairHmm.h:15,23,31), so the unmodified Java class
 * MicrosoftPairHmm binds to this library.
 *
 * Behaviour kept from ...MicrosoftPairHmm.cc:52-235: initNative(use_double, max_threads) records the precision mode
 * (max_threads has no meaning on a GPU); computeLikelihoodsNative takes the four direct buffers (IntBuffer lengths,
 * ByteBuffer reads, ByteBuffer haplotypes, DoubleBuffer likelihoods) and fills likelihoods[read * hapCount + hap].
 * The reference has no error channel; on a failure (no GPU, CUDA error) this writes nothing and prints the reason to
 * stderr -- it never computes on the CPU.  The GPU context is created by the first compute call.
 */
#include <jni.h>

#include <cstdio>
#include <mutex>

#include "../../include/mgl_pairhmm.h"

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
    {
        std::lock_guard<std::mutex> lk(g_mu);
        if (!g_handle) {
            const int rc = mgl_pairhmm_create(0, &g_handle);
            if (rc != MGL_PH_OK) { g_handle = nullptr; std::fprintf(stderr, "mgl_pairhmm computeLikelihoodsNative: no GPU context (code %d); no result written\n", rc); return; }
        }
    }
    const int32_t *lengths = (const int32_t *)env->GetDirectBufferAddress(lengthBuffer);
    const uint8_t *reads = (const uint8_t *)env->GetDirectBufferAddress(readsBuffer);
    const uint8_t *haps = (const uint8_t *)env->GetDirectBufferAddress(haplotypesBuffer);
    double *out = (double *)env->GetDirectBufferAddress(likelihoodBuffer);
    const int rc = mgl_pairhmm_compute(g_handle, lengths, reads, haps, g_use_double ? 1 : 0, out, nullptr);
    if (rc != MGL_PH_OK)
        std::fprintf(stderr, "mgl_pairhmm computeLikelihoodsNative: GPU computation failed (code %d: %s); no result written\n", rc,
                     mgl_pairhmm_last_error(g_handle));
}

JNIEXPORT void JNICALL Java_com_microsoft_mgl_pairhmm_MicrosoftPairHmm_doneNative(JNIEnv *, jobject) {}

}  // extern "C"

/*
 * sw_jni.cpp -- the three JNI symbols of the reference's libmgl_sw.so, with the
 * reference's exact signatures (src/main/native/mgl_sw/
 * com_microsoft_mgl_smithwaterman_MicrosoftSmithWaterman.h:15,23,31; JNI
 * signature (Ljava/nio/ByteBuffer;Ljava/nio/ByteBuffer;IIIIIII)I, .h:21), so the
 * unmodified Java class MicrosoftSmithWaterman.java:94-96 binds to this library.
 *
 * Behaviour kept from ...SmithWaterman.cpp:44-71: readsBuffer holds ref || alt,
 * scores are taken with any sign, the CIGAR text is written to cigarBuffer
 * WITHOUT a terminating NUL (cigar.copy, :65,:69 -- Java trims the zero-filled
 * buffer), the return value is the alignment offset.  The reference has no
 * error channel; on a failure (no GPU, CUDA error, over-long input) this writes
 * nothing and raises java.lang.RuntimeException (jni_throw.h) -- it never
 * silently computes on the CPU and never returns an empty result as if it were
 * one.  initNative has no caller in the Java class (.java:94), so the GPU
 * context is created lazily by the first alignNative; doneNative (called by
 * close(), .java:89-92) releases it, and a later call creates it again.
 *
 * Built against <jni.h> from a JDK when JAVA_HOME is set, else against the
 * ABI-faithful jni_min/jni.h in this directory.
 */
#include <jni.h>

#include <cstdio>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/mgl_sw.h"
#include "jni_throw.h"

extern "C" {

JNIEXPORT void JNICALL Java_com_microsoft_mgl_smithwaterman_MicrosoftSmithWaterman_initNative(JNIEnv *, jclass) {}

JNIEXPORT jint JNICALL Java_com_microsoft_mgl_smithwaterman_MicrosoftSmithWaterman_alignNative(
    JNIEnv *env, jclass, jobject readsBuffer, jobject cigarBuffer, jint targetLength, jint queryLength,
    jint match, jint mismatch, jint gapOpen, jint gapExtend, jint overhangStrategy)
{
    const uint8_t *target = (const uint8_t *)env->GetDirectBufferAddress(readsBuffer);
    char *cigar = (char *)env->GetDirectBufferAddress(cigarBuffer);
    if (!target || !cigar) { mgl_jni_throw(env, "mgl_sw alignNative: readsBuffer and cigarBuffer must be direct ByteBuffers"); return 0; }
    const uint8_t *query = target + targetLength;
    /* capacity the Java side allocates: 2 * max(refLength, altLength) (.java:71) */
    const int cap = 2 * (targetLength > queryLength ? targetLength : queryLength);
    std::vector<char> text((size_t)(targetLength + queryLength) * 12 + 16);
    int32_t len = 0, offset = 0;
    const int rc = mgl_sw_align_one(target, targetLength, query, queryLength, match, mismatch, gapOpen, gapExtend,
                                    overhangStrategy, text.data(), (int32_t)text.size(), &len, &offset);
    if (rc != MGL_SW_OK) {
        const std::string msg = "mgl_sw alignNative: GPU alignment failed (code " + std::to_string(rc) + "); no result written";
        mgl_jni_throw(env, msg.c_str());
        return 0;
    }
    /* the reference does not check capacity either (:65); copy what it would have copied */
    (void)cap;
    for (int k = 0; k < len; k++) cigar[k] = text[(size_t)k];
    return offset;
}

/* close() -> doneNative (.java:89-92): the process-wide handle and its device buffers go; the next call recreates them */
JNIEXPORT void JNICALL Java_com_microsoft_mgl_smithwaterman_MicrosoftSmithWaterman_doneNative(JNIEnv *, jclass) { mgl_sw_shutdown(); }

/*
 * The batched crossing (no counterpart in the reference, whose align() is one pair per call,
 * MicrosoftSmithWaterman.java:66-86).  Binds to
 *   private static native int alignBatchNative(ByteBuffer arena, long arenaBytes, int n,
 *           ByteBuffer refOff, ByteBuffer refLen, ByteBuffer altOff, ByteBuffer altLen,           // long[n], int[n], long[n], int[n]
 *           int match, int mismatch, int gapOpen, int gapExtend, int overhangStrategy, ByteBuffer strategies,  // int[n] or null
 *           ByteBuffer offsetsOut, ByteBuffer cigarStartOut, ByteBuffer cigarTextOut, long textCap);  // int[n], long[n+1], byte[textCap]
 * (INTEGRATION.md has the Java side).  All buffers are direct and in native byte order; pair p's CIGAR is the
 * characters cigarTextOut[cigarStartOut[p] .. cigarStartOut[p+1]), exactly what alignNative would have written for
 * it, formatted on the GPU.  Returns 0 or an MGL_SW_E_* code; on MGL_SW_E_OPS_CAPACITY cigarStartOut[n] holds the
 * text capacity needed (the Java side retries with a larger buffer); any other failure also raises
 * java.lang.RuntimeException.  Runs on the process-wide handle (mgl_sw_default_handle): every visible GPU, or the
 * MGL_SW_DEVICES list, the batch cut by DP cells across them.
 */
JNIEXPORT jint JNICALL Java_com_microsoft_mgl_smithwaterman_MicrosoftSmithWaterman_alignBatchNative(
    JNIEnv *env, jclass, jobject arena, jlong arenaBytes, jint n, jobject refOff, jobject refLen, jobject altOff, jobject altLen,
    jint match, jint mismatch, jint gapOpen, jint gapExtend, jint overhangStrategy, jobject strategies,
    jobject offsetsOut, jobject cigarStartOut, jobject cigarTextOut, jlong textCap)
{
    mgl_sw_handle *handle = nullptr;
    {
        const int rc = mgl_sw_default_handle(&handle);
        if (rc != MGL_SW_OK) { mgl_jni_throw(env, "mgl_sw alignBatchNative: no usable B200 (sm_100a) device"); return rc; }
    }
    if (n < 0 || arenaBytes < 0 || textCap < 0) { mgl_sw_default_release(); return MGL_SW_E_INVALID; }
    mgl_sw_batch b;
    b.n = n;
    b.arena = (const uint8_t *)env->GetDirectBufferAddress(arena);
    b.arena_bytes = (uint64_t)arenaBytes;
    b.ref_off = (const uint64_t *)env->GetDirectBufferAddress(refOff);
    b.ref_len = (const uint32_t *)env->GetDirectBufferAddress(refLen);
    b.alt_off = (const uint64_t *)env->GetDirectBufferAddress(altOff);
    b.alt_len = (const uint32_t *)env->GetDirectBufferAddress(altLen);
    b.strategies = strategies ? (const int32_t *)env->GetDirectBufferAddress(strategies) : nullptr;
    b.strategy = overhangStrategy;
    b.match = match; b.mismatch = mismatch; b.gap_open = gapOpen; b.gap_extend = gapExtend;
    mgl_sw_text_result r;
    r.offsets = (int32_t *)env->GetDirectBufferAddress(offsetsOut);
    r.text_start = (uint64_t *)env->GetDirectBufferAddress(cigarStartOut);
    r.text = (char *)env->GetDirectBufferAddress(cigarTextOut);
    r.text_cap = (uint64_t)textCap;
    r.text_total = 0;
    if (n > 0 && (!b.arena || !b.ref_off || !b.ref_len || !b.alt_off || !b.alt_len || !r.offsets || !r.text_start || (!r.text && textCap))) {
        mgl_sw_default_release();
        mgl_jni_throw(env, "mgl_sw alignBatchNative: every buffer must be a direct ByteBuffer");
        return MGL_SW_E_INVALID;
    }
    const int rc = mgl_sw_align_batch_text(handle, &b, &r, nullptr);
    if (rc == MGL_SW_E_OPS_CAPACITY && r.text_start) r.text_start[n] = r.text_total;
    else if (rc != MGL_SW_OK) {
        const std::string msg = std::string("mgl_sw alignBatchNative: failed (code ") + std::to_string(rc) + "): " + mgl_sw_last_error(handle);
        mgl_jni_throw(env, msg.c_str());
    }
    mgl_sw_default_release();
    return rc;
}

}  // extern "C"

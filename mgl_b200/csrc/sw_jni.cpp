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
 * error channel; on a failure (no GPU, CUDA error) this writes nothing and
 * returns 0 after printing the reason to stderr -- it never silently computes
 * on the CPU.  initNative has no caller in the Java class (.java:94), so the
 * GPU context is created lazily by the first alignNative.
 *
 * Built against <jni.h> from a JDK when JAVA_HOME is set, else against the
 * ABI-faithful jni_min/jni.h in this directory.
 */
#include <jni.h>

#include <cstdio>
#include <vector>

#include "../../include/mgl_sw.h"

extern "C" {

JNIEXPORT void JNICALL Java_com_microsoft_mgl_smithwaterman_MicrosoftSmithWaterman_initNative(JNIEnv *, jclass) {}

JNIEXPORT jint JNICALL Java_com_microsoft_mgl_smithwaterman_MicrosoftSmithWaterman_alignNative(
    JNIEnv *env, jclass, jobject readsBuffer, jobject cigarBuffer, jint targetLength, jint queryLength,
    jint match, jint mismatch, jint gapOpen, jint gapExtend, jint overhangStrategy)
{
    const uint8_t *target = (const uint8_t *)env->GetDirectBufferAddress(readsBuffer);
    char *cigar = (char *)env->GetDirectBufferAddress(cigarBuffer);
    if (!target || !cigar) return 0;
    const uint8_t *query = target + targetLength;
    /* capacity the Java side allocates: 2 * max(refLength, altLength) (.java:71) */
    const int cap = 2 * (targetLength > queryLength ? targetLength : queryLength);
    std::vector<char> text((size_t)(targetLength + queryLength) * 12 + 16);
    int32_t len = 0, offset = 0;
    const int rc = mgl_sw_align_one(target, targetLength, query, queryLength, match, mismatch, gapOpen, gapExtend,
                                    overhangStrategy, text.data(), (int32_t)text.size(), &len, &offset);
    if (rc != MGL_SW_OK) {
        std::fprintf(stderr, "mgl_sw alignNative: GPU alignment failed (code %d); no result written\n", rc);
        return 0;
    }
    /* the reference does not check capacity either (:65); copy what it would have copied */
    (void)cap;
    for (int k = 0; k < len; k++) cigar[k] = text[(size_t)k];
    return offset;
}

JNIEXPORT void JNICALL Java_com_microsoft_mgl_smithwaterman_MicrosoftSmithWaterman_doneNative(JNIEnv *, jclass) {}

}  // extern "C"

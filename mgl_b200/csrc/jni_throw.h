/*
 * jni_throw.h -- how a failed native call reaches Java.  The reference has no error channel at all
 * (...MicrosoftSmithWaterman.cpp:44-71 and ...MicrosoftPairHmm.cc:77-235 return whatever they computed); a GPU
 * library can fail (no device, CUDA error, over-long input), and returning as if nothing had happened would hand
 * GATK an empty CIGAR at offset 0 or a likelihood of 10^0.  So a failure raises java.lang.RuntimeException with
 * the reason, and the Java caller sees it at the call site.  Works with a JDK's <jni.h> and with jni_min/jni.h.
 * A harness whose JNIEnv table lacks FindClass / ThrowNew (the ctypes tables of tests/) gets the message on stderr.
 */
#ifndef MGL_B200_JNI_THROW_H
#define MGL_B200_JNI_THROW_H
#include <jni.h>

#include <cstdio>

static inline void mgl_jni_throw(JNIEnv *env, const char *msg)
{
    if (env && env->functions && env->functions->FindClass && env->functions->ThrowNew) {
        jclass cls = env->FindClass("java/lang/RuntimeException");
        if (cls && env->ThrowNew(cls, msg) == 0) return;
    }
    std::fprintf(stderr, "%s\n", msg);
}
#endif

/*
 * Minimal, ABI-faithful <jni.h> for building the JNI entry points on a machine
 * without a JDK (this image has none).  Only what the Smith-Waterman binding
 * touches is declared: the primitive typedefs, JNIEnv::GetDirectBufferAddress, and
 * FindClass / ThrowNew (a failed native call raises a Java RuntimeException).
 *
 * The JNIEnv layout follows the JNI specification's "Interface Function Table":
 * JNIEnv points to a table of function pointers in which slots 0-3 are reserved,
 * FindClass is slot 6, ThrowNew slot 14 and GetDirectBufferAddress slot 230.  A library compiled against this
 * header therefore calls the right slot of a real JVM's table.  When a JDK is
 * available, build with -I$JAVA_HOME/include -I$JAVA_HOME/include/linux instead
 * (mgl_b200/csrc/Makefile picks that up from JAVA_HOME).
 */
#ifndef MGL_B200_JNI_MIN_H
#define MGL_B200_JNI_MIN_H

#include <stdint.h>

#define JNIEXPORT __attribute__((visibility("default")))
#define JNIIMPORT
#define JNICALL

typedef int32_t jint;
typedef int64_t jlong;
typedef int8_t  jbyte;
typedef uint8_t jboolean;
typedef double  jdouble;

#ifdef __cplusplus
class _jobject {};
class _jclass : public _jobject {};
typedef _jobject *jobject;
typedef _jclass  *jclass;
#else
struct _jobject;
typedef struct _jobject *jobject;
typedef jobject jclass;
#endif

#define MGL_JNI_SLOT_FindClass 6
#define MGL_JNI_SLOT_ThrowNew 14
#define MGL_JNI_SLOT_GetDirectBufferAddress 230
#define MGL_JNI_TABLE_SLOTS 235

struct JNINativeInterface_;
struct JNIEnv_;
#ifdef __cplusplus
typedef JNIEnv_ JNIEnv;
#else
typedef const struct JNINativeInterface_ *JNIEnv;
#endif

struct JNINativeInterface_ {
    void *slots_0_5[MGL_JNI_SLOT_FindClass];
    jclass (JNICALL *FindClass)(JNIEnv *env, const char *name);
    void *slots_7_13[MGL_JNI_SLOT_ThrowNew - MGL_JNI_SLOT_FindClass - 1];
    jint (JNICALL *ThrowNew)(JNIEnv *env, jclass clazz, const char *msg);
    void *slots_15_229[MGL_JNI_SLOT_GetDirectBufferAddress - MGL_JNI_SLOT_ThrowNew - 1];
    void *(JNICALL *GetDirectBufferAddress)(JNIEnv *env, jobject buf);
    void *slots_231_234[MGL_JNI_TABLE_SLOTS - MGL_JNI_SLOT_GetDirectBufferAddress - 1];
};

#ifdef __cplusplus
struct JNIEnv_ {
    const struct JNINativeInterface_ *functions;
    void *GetDirectBufferAddress(jobject buf) { return functions->GetDirectBufferAddress(this, buf); }
    jclass FindClass(const char *name) { return functions->FindClass(this, name); }
    jint ThrowNew(jclass clazz, const char *msg) { return functions->ThrowNew(this, clazz, msg); }
};
#endif

#endif /* MGL_B200_JNI_MIN_H */

/* Minimal stand-in for <jni.h>, ONLY so that tests can check that integration/.../fmlk_jni.cc parses
 * and type-checks against include/fmlk.h where no JDK is installed.  It declares the handful of JNI
 * types and JNIEnv members the shim uses, with the signatures of the JNI specification; nothing
 * here is ever linked or run.  A real build uses the JDK's jni.h. */
#ifndef FMLK_CHECK_JNI_H
#define FMLK_CHECK_JNI_H
#include <cstdint>

#define JNIEXPORT
#define JNICALL
typedef int32_t jint;
typedef int64_t jlong;
typedef int8_t jbyte;
typedef double jdouble;
typedef jint jsize;
typedef uint8_t jboolean;
class _jobject {};
typedef _jobject *jobject;
typedef jobject jclass;
typedef jobject jstring;
typedef jobject jarray;
typedef jarray jbyteArray;
typedef jarray jintArray;
typedef jarray jlongArray;
typedef jarray jdoubleArray;

struct JNIEnv {
    jclass FindClass(const char *name);
    jint ThrowNew(jclass cls, const char *msg);
    jstring NewStringUTF(const char *utf);
    void *GetDirectBufferAddress(jobject buf);
    jsize GetArrayLength(jarray a);
    jbyteArray NewByteArray(jsize n);
    void SetByteArrayRegion(jbyteArray a, jsize start, jsize n, const jbyte *buf);
    void GetByteArrayRegion(jbyteArray a, jsize start, jsize n, jbyte *buf);
    void SetIntArrayRegion(jintArray a, jsize start, jsize n, const jint *buf);
    void SetLongArrayRegion(jlongArray a, jsize start, jsize n, const jlong *buf);
    void GetLongArrayRegion(jlongArray a, jsize start, jsize n, jlong *buf);
    jdouble *GetDoubleArrayElements(jdoubleArray a, jboolean *is_copy);
    void ReleaseDoubleArrayElements(jdoubleArray a, jdouble *elems, jint mode);
};
#endif

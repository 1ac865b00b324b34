/* jni.h STUB — test infrastructure only (tests/test_binding_cpu.py): just enough of the JNI surface that
 * jni/hovl_jni.c uses for `gcc -fsyntax-only` to type-check it on a box without a JDK.  Signatures follow the
 * JNI specification; nothing here is ever linked or run. */
#ifndef HOVL_JNI_STUB_H
#define HOVL_JNI_STUB_H
#include <stdint.h>
typedef int32_t jint; typedef int64_t jlong; typedef int8_t jbyte; typedef uint8_t jboolean; typedef jint jsize;
typedef void *jobject; typedef jobject jclass; typedef jobject jarray; typedef jarray jintArray; typedef jarray jlongArray;
typedef jarray jbyteArray;
#define JNIEXPORT __attribute__((visibility("default")))
#define JNICALL
#define JNI_ABORT 2
struct JNINativeInterface_;
typedef const struct JNINativeInterface_ *JNIEnv;
struct JNINativeInterface_ {
  jclass (*FindClass)(JNIEnv *, const char *);
  jint (*ThrowNew)(JNIEnv *, jclass, const char *);
  jsize (*GetArrayLength)(JNIEnv *, jarray);
  void (*GetIntArrayRegion)(JNIEnv *, jintArray, jsize, jsize, jint *);
  void (*SetIntArrayRegion)(JNIEnv *, jintArray, jsize, jsize, const jint *);
  void (*SetByteArrayRegion)(JNIEnv *, jbyteArray, jsize, jsize, const jbyte *);
  jintArray (*NewIntArray)(JNIEnv *, jsize);
  jbyteArray (*NewByteArray)(JNIEnv *, jsize);
  void *(*GetPrimitiveArrayCritical)(JNIEnv *, jarray, jboolean *);
  void (*ReleasePrimitiveArrayCritical)(JNIEnv *, jarray, void *, jint);
  void *(*GetDirectBufferAddress)(JNIEnv *, jobject);
};
#endif

/* hovl_jni.c — JNI shim between the reference's Scala host and libhovl.so (include/hovl.h).
 *
 * The reference (rohit507/Sequence-Aligner) has no FFI; this file is what a maintainer adds so that
 * Project4.main (src/Project4.scala:56-60) can replace
 *     generateKmerTable(input)                    src/Project4.scala:508
 *     genAlignment(table, alignSettings, true)    src/Project4.scala:567
 * with one native call.  readArgs / readSeq / readHOXD / calcOverlaps stay on the JVM (host/ directory).
 *
 * Build (needs a JDK; none exists in the build image, tests/test_binding_cpu.py only syntax-checks this file
 * against a stub jni.h):
 *   gcc -shared -fPIC -I$JAVA_HOME/include -I$JAVA_HOME/include/linux -Iinclude jni/hovl_jni.c \
 *       -Lsequence-aligner_b200 -lhovl -Wl,-rpath,'$ORIGIN' -o sequence-aligner_b200/libhovl_jni.so
 *
 * Ownership: `bases` (a direct ByteBuffer) and `offsets` are borrowed for the duration of the call; the records
 * are copied into a fresh int[] and the native result is freed before returning.  Errors become
 * RuntimeException(hovl_last_error) — the reference's own failure modes are unchecked exceptions too
 * (StringIndexOutOfBounds at src/BioLibs.scala:648 <-> HOVL_E_SHORT_READ, MatchError at :144 <-> HOVL_E_BAD_BASE). */
#include <jni.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "hovl.h"

#define HOVL_SETTINGS_INTS ((jsize)(sizeof(hovl_settings) / 4))   /* 32: every field is 32 bits wide */
#define HOVL_RECORD_INTS   ((jsize)(sizeof(hovl_record) / 4))     /* 13 */

static void throw_rt(JNIEnv *e, const char *msg) {
  jclass c = (*e)->FindClass(e, "java/lang/RuntimeException");
  if (c) (*e)->ThrowNew(e, c, msg ? msg : "hovl: unknown error");
}

/* long HovlNative.create(int device) */
JNIEXPORT jlong JNICALL Java_HovlNative_create(JNIEnv *e, jclass cls, jint device) {
  hovl_ctx *ctx = NULL;
  int dev = (int)device;
  (void)cls;
  if (hovl_create(&ctx, &dev, 1) != HOVL_OK) { throw_rt(e, hovl_last_error(NULL)); return 0; }
  return (jlong)(intptr_t)ctx;
}

/* long HovlNative.createMulti(int[] devices): ONE handle over several GPUs of this process (include/hovl.h, hovl_create
 * with n_devices > 1; csrc/multi.cu).  run() on it shards the reads, exchanges k-mers and partial pair counts over peer
 * memory and returns the records in canonical (a, b) order: it needs order = 1 and id_mode = 1 in the settings
 * (NativeOverlap.pack(..., wide = true)).  No Python, NCCL or launcher on the JVM side. */
JNIEXPORT jlong JNICALL Java_HovlNative_createMulti(JNIEnv *e, jclass cls, jintArray jdev) {
  hovl_ctx *ctx = NULL;
  jint dev[64];
  int d[64];
  jsize n, i;
  (void)cls;
  if (!jdev) { throw_rt(e, "hovl: null argument"); return 0; }
  n = (*e)->GetArrayLength(e, jdev);
  if (n < 1 || n > 64) { throw_rt(e, "hovl: 1..64 devices"); return 0; }
  (*e)->GetIntArrayRegion(e, jdev, 0, n, dev);
  for (i = 0; i < n; i++) d[i] = (int)dev[i];
  if (hovl_create(&ctx, d, (int)n) != HOVL_OK) { throw_rt(e, hovl_last_error(NULL)); return 0; }
  return (jlong)(intptr_t)ctx;
}

/* void HovlNative.destroy(long ctx) */
JNIEXPORT void JNICALL Java_HovlNative_destroy(JNIEnv *e, jclass cls, jlong h) {
  (void)e; (void)cls;
  hovl_destroy((hovl_ctx *)(intptr_t)h);
}

/* int[] HovlNative.run(long ctx, int[] settings, java.nio.ByteBuffer bases, long[] offsets)
 *   settings : hovl_settings as HOVL_SETTINGS_INTS ints (host/NativeOverlap.scala packs them field by field)
 *   bases    : DIRECT buffer, upper-case ASCII of all reads concatenated
 *   offsets  : n_reads + 1 entries; read i (id i+1, src/BioLibs.scala:27) = bases[offsets[i] .. offsets[i+1])
 * returns the records flattened, HOVL_RECORD_INTS ints per record, in the reference's dispatch order. */
JNIEXPORT jintArray JNICALL Java_HovlNative_run(JNIEnv *e, jclass cls, jlong h, jintArray jset, jobject jbases,
                                                jlongArray joffs) {
  hovl_ctx *ctx = (hovl_ctx *)(intptr_t)h;
  hovl_settings s;
  hovl_result *res = NULL;
  const uint8_t *bases;
  jlong *offs;
  jsize n1;
  int rc;
  uint64_t n = 0;
  const hovl_record *recs;
  jintArray out;
  (void)cls;
  if (!ctx || !jset || !joffs) { throw_rt(e, "hovl: null argument"); return NULL; }
  if ((*e)->GetArrayLength(e, jset) != HOVL_SETTINGS_INTS) { throw_rt(e, "hovl: settings must hold sizeof(hovl_settings)/4 ints"); return NULL; }
  (*e)->GetIntArrayRegion(e, jset, 0, HOVL_SETTINGS_INTS, (jint *)&s);
  n1 = (*e)->GetArrayLength(e, joffs);
  if (n1 < 1) { throw_rt(e, "hovl: offsets must hold n_reads + 1 entries"); return NULL; }
  bases = jbases ? (const uint8_t *)(*e)->GetDirectBufferAddress(e, jbases) : NULL;
  offs = (jlong *)(*e)->GetPrimitiveArrayCritical(e, joffs, NULL);
  if (!offs) { throw_rt(e, "hovl: cannot pin offsets"); return NULL; }
  rc = hovl_run(ctx, &s, bases, (const uint64_t *)offs, (uint32_t)(n1 - 1), &res);
  (*e)->ReleasePrimitiveArrayCritical(e, joffs, offs, JNI_ABORT);
  if (rc != HOVL_OK) {
    if (res) hovl_result_free(res);
    throw_rt(e, hovl_last_error(ctx));
    return NULL;
  }
  recs = hovl_records(res, &n);
  if (n > (uint64_t)(0x7fffffff / HOVL_RECORD_INTS)) { hovl_result_free(res); throw_rt(e, "hovl: more records than a Java array holds"); return NULL; }
  out = (*e)->NewIntArray(e, (jsize)n * HOVL_RECORD_INTS);
  if (out && n) (*e)->SetIntArrayRegion(e, out, 0, (jsize)n * HOVL_RECORD_INTS, (const jint *)recs);
  hovl_result_free(res);
  return out;
}

/* byte[] HovlNative.formatOvl(int[] records): Overlap.print + calcOverlaps text (src/ObjectStore.scala:127-135,
 * src/Project4.scala:806-820) for hosts that want the native writer; the JVM writer works unchanged as well. */
JNIEXPORT jbyteArray JNICALL Java_HovlNative_formatOvl(JNIEnv *e, jclass cls, jintArray jrecs) {
  jsize len;
  jint *r;
  uint64_t need;
  jbyteArray out;
  char *buf;
  (void)cls;
  if (!jrecs) { throw_rt(e, "hovl: null argument"); return NULL; }
  len = (*e)->GetArrayLength(e, jrecs);
  r = (jint *)(*e)->GetPrimitiveArrayCritical(e, jrecs, NULL);
  if (!r) { throw_rt(e, "hovl: cannot pin records"); return NULL; }
  need = hovl_format_ovl((const hovl_record *)r, (uint64_t)(len / HOVL_RECORD_INTS), NULL, 0);
  buf = (char *)malloc(need ? need : 1);
  if (buf) hovl_format_ovl((const hovl_record *)r, (uint64_t)(len / HOVL_RECORD_INTS), buf, need);
  (*e)->ReleasePrimitiveArrayCritical(e, jrecs, r, JNI_ABORT);
  if (!buf || need > 0x7fffffffu) { free(buf); throw_rt(e, "hovl: out of memory"); return NULL; }
  out = (*e)->NewByteArray(e, (jsize)need);
  if (out && need) (*e)->SetByteArrayRegion(e, out, 0, (jsize)need, (const jbyte *)buf);
  free(buf);
  return out;
}

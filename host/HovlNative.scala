// HovlNative.scala — the JNI surface of libhovl_jni.so (jni/hovl_jni.c) over libhovl.so (include/hovl.h).
// Scala 2.9 (the reference's compiler, Rakefile.rb:66-71).  Not compiled in the build image (no JVM there).
object HovlNative {
  System.loadLibrary("hovl_jni")
  @native def create(device: Int): Long
  /** One handle over several GPUs of this process; run() then needs order = 1, id_mode = 1 (canonical record order). */
  @native def createMulti(devices: Array[Int]): Long
  @native def destroy(ctx: Long): Unit
  /** settings: hovl_settings as 32 ints; bases: DIRECT buffer; offsets: n_reads + 1 entries.
    * Returns 13 ints per record (hovl_record), in the reference's dispatch order. */
  @native def run(ctx: Long, settings: Array[Int], bases: java.nio.ByteBuffer, offsets: Array[Long]): Array[Int]
  @native def formatOvl(records: Array[Int]): Array[Byte]
}

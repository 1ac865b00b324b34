// NativeOverlap.scala — drop-in for the two calls Project4.main makes into the hash-overlap path
//     generateKmerTable(input)                    src/Project4.scala:508
//     genAlignment(table, alignSettings, filter)  src/Project4.scala:567
// readArgs, readSeq, readHOXD, calcOverlaps, Alignment, Overlap stay exactly as they are in the reference.
// Scala 2.9.  Not compiled in the build image (no JVM there); checked by reading against the reference:
//   Alignment(sA, sB, oA: String, oB: String, st, en, cor, err)      src/ObjectStore.scala:88-89
//   Alignment.valid reads alignA.length (>= minOverlap)              src/ObjectStore.scala:102-107
//   AlignSettings.costFunc is a (Char, Char) => Int                  src/ObjectStore.scala:16-19
//   calcOverlaps(alignments, settings) writes to Project4.output     src/Project4.scala:795-825
object NativeOverlap {
  import java.lang.Float.floatToRawIntBits

  val Bases = "ACGT"                       // row / column order of hovl_settings.cost (include/hovl.h)

  /** AlignSettings (src/ObjectStore.scala:16-36) -> hovl_settings, field by field, 32 ints.
    * The float thresholds cross as raw bits, so they are the JVM's own float32 values. */
  def pack(s: AlignSettings, quadratic: Boolean, filter: Boolean, wide: Boolean = false): Array[Int] = {
    val a = new Array[Int](32)
    a(0) = s.kmerSize
    for (i <- 0 until 4; j <- 0 until 4) a(1 + 4 * i + j) = s.costFunc(Bases(i), Bases(j))   // BioLibs.scala:119-160
    a(17) = s.gapOpen; a(18) = s.gapExtend; a(19) = s.minOverlap
    a(20) = floatToRawIntBits(s.minIdentity); a(21) = floatToRawIntBits(s.maxIgnore)
    a(22) = s.minCollisions; a(23) = s.maxCollisions
    a(24) = floatToRawIntBits(s.kmerHeadEdge);    a(25) = floatToRawIntBits(s.kmerTailEdge)
    a(26) = floatToRawIntBits(s.kmerMidLeadEdge); a(27) = floatToRawIntBits(s.kmerMidTailEdge)
    a(28) = if (quadratic) 1 else 0        // --quadratic-align / --linear-align, src/Project4.scala:187-192
    a(29) = if (wide) 1 else 0             // order: 0 = the reference's (Trove iteration) order, 1 = canonical (a, b)
    a(30) = if (wide) 1 else 0             // id_mode: 0 = the reference's 16-bit pair key (<= 32767 reads), 1 = wide key
    a(31) = if (filter) 1 else 0           // genAlignment's `filter`
    a
  }

  /** generateKmerTable + genAlignment in one native call.  More than one device, or more reads than the reference's
    * 16-bit pair key holds (src/KmerTable.scala:73), switch to the wide key and canonical record order: tigger does not
    * depend on the order of the OVL records (tests/test_amos_pipeline_cpu.py). */
  def align(file: String, s: AlignSettings, filter: Boolean, quadratic: Boolean,
            devices: Array[Int] = Array(0)): Seq[Alignment] = {
    val seqs = new scala.collection.mutable.ArrayBuffer[Sequence]
    BioLibs.readSeq(file, (q: Sequence) => seqs += q)                 // the reference's reader, src/BioLibs.scala:26-50
    val offs = new Array[Long](seqs.length + 1)
    for (i <- 0 until seqs.length) offs(i + 1) = offs(i) + seqs(i).seq.length
    val buf = java.nio.ByteBuffer.allocateDirect(math.max(offs(seqs.length).toInt, 1))
    for (q <- seqs) buf.put(q.seq.getBytes("US-ASCII"))
    val wide = devices.length > 1 || seqs.length > 32767
    val ctx = if (devices.length == 1) HovlNative.create(devices(0)) else HovlNative.createMulti(devices)
    val r = try { HovlNative.run(ctx, pack(s, quadratic, filter, wide), buf, offs) } finally { HovlNative.destroy(ctx) }
    val out = new scala.collection.mutable.ArrayBuffer[Alignment](r.length / 13)
    var o = 0
    while (o < r.length) {                                            // hovl_record: a b count flags si sj ei ej c e score ahg bhg
      val c = r(o + 8); val e = r(o + 9)
      // Alignment.valid only needs alignA.length == correct + error (every aligned column is a match or an error,
      // src/BioLibs.scala:541-575); the gapped strings themselves are not used downstream of valid().
      val pad = "-" * (c + e)
      if ((r(o + 3) & 1) != 0) out += BioLibs.dud                     // HOVL_F_DUD (only with filter = false)
      else out += new Alignment(seqs(r(o) - 1), seqs(r(o + 1) - 1), pad, pad,
                                (r(o + 4), r(o + 5)), (r(o + 6), r(o + 7)), c, e)
      o += 13
    }
    out
  }
}

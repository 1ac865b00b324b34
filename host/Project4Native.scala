// Project4Native.scala — Project4 with the hash-overlap path on the GPU.  Same command line, same output.
// Everything except the two calls below is the reference's own code: readArgs (src/Project4.scala:101-259),
// calcOverlaps (:795-825), BioLibs.readSeq / readHOXD, Alignment / Overlap (src/ObjectStore.scala:88-142).
// Compile with the reference's sources on the classpath (Rakefile.rb:66-71 order, this file last):
//   scalac -optimize -d build src/ObjectStore.scala src/BioLibs.scala src/KmerTable.scala src/Project4.scala \
//          host/HovlNative.scala host/NativeOverlap.scala host/Project4Native.scala
//   scala -cp build:lib/trove.jar -Djava.library.path=sequence-aligner_b200 Project4Native -i reads.seq -o out.ovl
object Project4Native {
  def main(args: Array[String]) {
    val settings = Project4.readArgs(args)                // sets Project4.input / output / action / fdAlign ...
    BioLibs.debug = Project4.debug
    Project4.action match {
      case "calc-overlaps" =>
        // was: var table = generateKmerTable(input); calcOverlaps(genAlignment(table, alignSettings, true), alignSettings)
        //      (src/Project4.scala:56-59)
        Project4.calcOverlaps(NativeOverlap.align(Project4.input, settings, true, !Project4.fdAlign), settings)
        System.exit(0)
      case _ =>
        Project4.main(args)                               // every other mode: the reference, unchanged
    }
  }
}

"""Host-side readers and the OVL writer: the parts of the reference that stay on the host
(BioLibs.readSeq / readHOXD, Project4.calcOverlaps), restated for the Python host mirror."""
import os

import numpy as np


def read_seq(path):
    """readSeq (src/BioLibs.scala:26-50): the first line must start with '>'; every later '>'
    line closes a record; header text is ignored and ids are 1,2,3.. in file order; body lines
    are concatenated and upper-cased; the last record is emitted at EOF; a line ends at "\n", "\r" or
    "\r\n" (BufferedReader.readLine).  Returns owned arrays (bases uint8[], offsets uint64[n+1]).

    One implementation for every host: this is hovl_reads_open (csrc/hostio.cpp), the reader hovl_cli uses, so a
    file is accepted or rejected the same way whichever host reads it.  Bad input raises ValueError."""
    from . import HovlError, Reads
    try:
        with Reads(path) as r:
            return np.array(r.bases, np.uint8), np.array(r.offsets, np.uint64)
    except HovlError as e:
        raise ValueError(str(e)) from None


def read_hoxd(path):
    """readHOXD (src/BioLibs.scala:66-114): skip the title line, read the comma-separated
    column header, then one row per base until EOF or an empty line.  Returns a 4x4 list in
    A,C,G,T order.  As in the reference the labels are trimmed and the numbers are NOT (Integer.parseInt(row(i)):
    " -114" is a NumberFormatException), a row longer than the header is an error (ArrayIndexOutOfBounds), and the
    pair-list form (amos/HOXD2.txt, 'A,C=-114') is rejected.  This is hovl_read_hoxd (csrc/hostio.cpp); bad input
    raises ValueError."""
    from . import HovlError, read_hoxd_native
    try:
        return read_hoxd_native(path)
    except HovlError as e:
        raise ValueError(str(e)) from None


def write_ovl(path, ovl_bytes):
    """calcOverlaps with -o (src/Project4.scala:808-820): an existing file is replaced."""
    if os.path.exists(path):
        os.remove(path)
    with open(path, "wb") as f:
        f.write(ovl_bytes)


def write_fasta(path, bases, offsets, prefix="r"):
    with open(path, "wb") as f:
        for i in range(len(offsets) - 1):
            f.write(b">%s%d\n" % (prefix.encode(), i + 1))
            f.write(bytes(bases[int(offsets[i]):int(offsets[i + 1])]))
            f.write(b"\n")

"""Host-side readers and the OVL writer: the parts of the reference that stay on the host
(BioLibs.readSeq / readHOXD, Project4.calcOverlaps), restated for the Python host mirror."""
import os

import numpy as np


def read_seq(path):
    """readSeq (src/BioLibs.scala:26-50): the first line must start with '>'; every later '>'
    line closes a record; header text is ignored and ids are 1,2,3.. in file order; body lines
    are concatenated and upper-cased; the last record is emitted at EOF.
    Returns (bases uint8[], offsets uint64[n+1])."""
    with open(path, "rb") as f:
        data = f.read()
    lines = data.split(b"\n")
    if lines and lines[-1] == b"":
        lines.pop()
    if not lines or not lines[0].startswith(b">"):
        raise ValueError("Invalid Sequence File: %s" % path)
    chunks, offsets, total = [], [0], 0
    for ln in lines[1:]:
        if ln.endswith(b"\r"):
            ln = ln[:-1]
        if ln.startswith(b">"):
            offsets.append(total)
        else:
            up = ln.upper()
            chunks.append(up)
            total += len(up)
    offsets.append(total)
    bases = np.frombuffer(b"".join(chunks), np.uint8).copy()
    return bases, np.asarray(offsets, np.uint64)


_ACGT = {"A": 0, "C": 1, "G": 2, "T": 3}


def read_hoxd(path):
    """readHOXD (src/BioLibs.scala:66-114): skip the title line, read the comma-separated
    column header, then one row per base until EOF or an empty line.  Returns a 4x4 list in
    A,C,G,T order.  The pair-list form (amos/HOXD2.txt, 'A,C=-114') is rejected exactly like
    the reference's Integer.parseInt would."""
    with open(path, "r") as f:
        lines = f.read().split("\n")
    if len(lines) < 2:
        raise ValueError("bad HOXD file: %s" % path)
    cols = [c.strip() for c in lines[1].split(",")]
    m = [[0] * 4 for _ in range(4)]
    for ln in lines[2:]:
        ln = ln.rstrip("\r")
        if ln == "":
            break
        row = ln.split(",")
        a = row[0].strip()[:1].upper()
        for i in range(1, min(len(row), len(cols))):
            b = cols[i][:1].upper()
            if a not in _ACGT or b not in _ACGT:
                raise ValueError("bad HOXD base label in %s" % path)          # MatchError
            m[_ACGT[a]][_ACGT[b]] = int(row[i].strip())                          # NumberFormatException
    return m


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

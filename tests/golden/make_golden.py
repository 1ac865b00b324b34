#!/usr/bin/env python
"""Regenerates tests/golden/ from the read-only reference checkout (/root/reference).

Run in the build container only (the GPU box has no /root/reference); everything it writes is
committed.  What it makes and why:

  crp177.seq / crp177.ovl / crp177.fasta / HOXD1.txt / HOXD2.txt
      the reference's own small fixtures (amos/small, amos/), copied verbatim.
  c_ruddii.fasta + c_ruddii_layout.npy
      the reference ships the c_ruddii CONTIG (amos/c_ruddii.fasta) and the AMOS bank
      (amos/c_ruddii.bnk) but not the reads (missing blob).  The bank's layout
      (LAY.0.0.var: 1 header byte, then 32000 x 6 little-endian int32
      {0, read iid, 0, offset, clr_begin=0, clr_end=100}) places every read on the contig, all
      forward, so read iid = contig[offset : offset+100].  The layout offsets are stored in
      iid order; tests rebuild the 32000 reads from them.
  c_ruddii_amos_ovl.npz / crp177 golden
      AMOS's own overlaps for those reads (the reference's amos/hash-overlap run on a bank made
      with amos/toAmos_new; 388,898 records = amos/c_ruddii.bnk/OVL.ifo), as (a, b, ahg, bhg).
      These are a SUPERSET pin: every overlap the Scala algorithm emits must be in them with the
      same hangs.
  *_oracle_*.npz / *.oracle.ovl
      outputs of the CPU oracle (oracle/hovl_oracle.c) on the fixtures, so that the GPU tests
      compare against fixed vectors as well as against the live oracle.
"""
import hashlib
import os
import re
import shutil
import struct
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "sequence-aligner_b200"))


def parse_ovl_text(text):
    recs = re.findall(r"rds:(\d+),(\d+)\s+scr:(-?\d+)\s+ahg:(-?\d+)\s+bhg:(-?\d+)", text)
    return np.array([[int(a), int(b), int(c), int(d)] for a, b, _, c, d in recs], np.int32).reshape(-1, 4)


def main():
    from oracle import oracle as O
    from hovl import io as hio
    for src, dst in (("amos/small/crp177.seq", "crp177.seq"), ("amos/small/crp177.ovl", "crp177.ovl"),
                     ("amos/small/crp177.fasta", "crp177.fasta"), ("amos/HOXD1.txt", "HOXD1.txt"),
                     ("amos/HOXD2.txt", "HOXD2.txt"), ("amos/c_ruddii.fasta", "c_ruddii.fasta")):
        shutil.copyfile(os.path.join(REF, src), os.path.join(HERE, dst))
        os.chmod(os.path.join(HERE, dst), 0o644)

    # ---- c_ruddii reads from the bank layout
    raw = open(os.path.join(REF, "amos/c_ruddii.bnk/LAY.0.0.var"), "rb").read()
    n = (len(raw) - 1) // 24
    tiles = np.frombuffer(raw[1:1 + 24 * n], "<i4").reshape(n, 6)
    assert n == 32000 and (tiles[:, 4] == 0).all() and (tiles[:, 5] == 100).all()
    offsets = np.zeros(n, np.int32)
    offsets[tiles[:, 1] - 1] = tiles[:, 3]
    np.save(os.path.join(HERE, "c_ruddii_layout.npy"), offsets)
    contig, co = hio.read_seq(os.path.join(HERE, "c_ruddii.fasta"))
    assert len(co) == 2
    idx = offsets[:, None].astype(np.int64) + np.arange(100)[None, :]
    bases = contig[idx].reshape(-1)
    roffs = (np.arange(n + 1) * 100).astype(np.uint64)

    tmp = tempfile.mkdtemp(prefix="golden_")
    amos = os.path.join(REF, "amos")

    def amos_overlaps(name, b, o):
        seq = os.path.join(tmp, name + ".seq")
        bnk = os.path.join(tmp, name + ".bnk")
        hio.write_fasta(seq, b, o)
        subprocess.check_call([os.path.join(amos, "toAmos_new"), "-s", seq, "-b", bnk], stdout=subprocess.DEVNULL,
                              stderr=subprocess.DEVNULL)
        subprocess.check_call([os.path.join(amos, "hash-overlap"), bnk, "-B", "-x", "0.04", "-o", "40"],
                              stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        txt = subprocess.run([os.path.join(amos, "bank-report"), "-b", bnk, "OVL"], stdout=subprocess.PIPE,
                             stderr=subprocess.DEVNULL, check=True).stdout.decode()
        return parse_ovl_text(txt)

    if "--skip-amos" not in sys.argv:
        ov = amos_overlaps("c_ruddii", bases, roffs)
        assert len(ov) == 388898, len(ov)
        np.savez_compressed(os.path.join(HERE, "c_ruddii_amos_ovl.npz"), ovl=ov)

    # ---- oracle vectors
    def save_oracle(tag, s, b, o, with_text):
        r = O.run(s, b, o, threads=8, keep_stages=True)
        ovl = O.format_ovl(r.records)
        extra = {}
        if with_text:                 # small sets: keep PairData itself
            extra = dict(pair_key=r.pair_key, pair_count=r.pair_count)
        np.savez_compressed(os.path.join(HERE, tag + ".npz"), records=r.records,
                            stats=np.array([r.stats[k] for k in O.STATS_FIELDS[:14]], np.int64),
                            pair_sha256=np.frombuffer(hashlib.sha256(
                                r.pair_key.tobytes() + r.pair_count.tobytes()).digest(), np.uint8),
                            ovl_sha256=np.frombuffer(hashlib.sha256(ovl).digest(), np.uint8), **extra)
        if with_text:
            open(os.path.join(HERE, tag + ".ovl"), "wb").write(ovl)
        print(tag, {k: r.stats[k] for k in O.STATS_FIELDS[:14]})

    cb, co2 = hio.read_seq(os.path.join(HERE, "crp177.seq"))
    save_oracle("crp177_oracle_k12", O.default_settings(), cb, co2, True)
    hoxd = hio.read_hoxd(os.path.join(HERE, "HOXD1.txt"))
    save_oracle("crp177_oracle_k16", O.default_settings(k=16, cost=hoxd), cb, co2, True)
    save_oracle("c_ruddii_oracle_k12", O.default_settings(), bases, roffs, False)
    shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()

"""hovl_cli on a GPU: the Project4 command line end to end (native FASTA ingest -> hovl_run -> OVL file)
against the committed golden bytes and the CPU oracle."""
import os
import subprocess

import numpy as np
import pytest

from conftest import GOLDEN, java_float as _java_float, want_test_overlaps

pytestmark = pytest.mark.gpu


def cli(*args):
    import hovl
    assert os.path.exists(hovl.CLI_PATH), "hovl_cli not built"
    return subprocess.run([hovl.CLI_PATH] + list(args), capture_output=True, timeout=300)


@pytest.mark.parametrize("tag,flags", [("k12", []), ("k16", ["-k", "16", "-m", os.path.join(GOLDEN, "HOXD1.txt")])])
def test_calc_overlaps_bytes(tmp_path, tag, flags):
    """`Project4 -i crp177.seq [-k 16 -m HOXD1.txt] -o out.ovl` (K1-shaped and K2): identical bytes."""
    want = open(os.path.join(GOLDEN, "crp177_oracle_%s.ovl" % tag), "rb").read()
    out = tmp_path / "out.ovl"
    out.write_bytes(b"stale")
    p = cli("-i", os.path.join(GOLDEN, "crp177.seq"), "-o", str(out), "--hovl-stats", *flags)
    assert p.returncode == 0, p.stderr
    assert out.read_bytes() == want
    assert b"kernel launches" in p.stderr
    p = cli("-i", os.path.join(GOLDEN, "crp177.seq"), "--mt-hash", "--block-align", "--linear-align", *flags)   # stdout form
    assert p.returncode == 0 and p.stdout == want


def test_dispatch_probe_modes(oracle, crp177):
    """--test-dispatch-collisions / --test-block-dispatch (src/Project4.scala:376-421) from the native records."""
    bases, offs = crp177
    s = oracle.default_settings(filter_valid=0)
    ref = oracle.run(s, bases, offs, threads=4).records
    lines = [" Dispatched Coll : %d - %d <-> %d" % (i + 1, r["a"], r["b"]) for i, r in enumerate(ref)]
    p = cli("-i", os.path.join(GOLDEN, "crp177.seq"), "--test-dispatch-collisions")
    assert p.returncode == 0, p.stderr
    assert p.stdout.decode().split("\n")[:-1] == lines
    a = ref["a"]
    runs = np.diff(np.flatnonzero(np.concatenate(([True], a[1:] != a[:-1], [True]))))
    hist = {}
    for n in runs:
        hist[int(n)] = hist.get(int(n), 0) + 1
    tail = ["", " Histogram Of Relations : [Number of Aligns -> Number of Seqs w/ that many Aligns]"] + \
           ["          [%d -> %d]" % (k, hist[k]) for k in sorted(hist)] + ["", ""]
    p = cli("-i", os.path.join(GOLDEN, "crp177.seq"), "--test-block-dispatch")
    assert p.returncode == 0, p.stderr
    assert p.stdout.decode().split("\n") == lines + tail


def test_reference_id_limit_and_wide_ids(tmp_path, oracle):
    """> 32767 reads collide in the reference's 16-bit pair key (src/KmerTable.scala:73): refused unless --hovl-wide-ids."""
    from conftest import synth_reads
    from hovl import io as hio
    bases, offs = synth_reads(5, 33000, 60, 40000)
    fa = tmp_path / "many.fa"
    hio.write_fasta(str(fa), bases, offs)
    p = cli("-i", str(fa))
    assert p.returncode == 1 and b"error -2" in p.stderr
    out = tmp_path / "wide.ovl"
    p = cli("-i", str(fa), "--hovl-wide-ids", "-o", str(out))
    assert p.returncode == 0, p.stderr
    ref = oracle.run(oracle.default_settings(order=1, id_mode=1, filter_valid=1), bases, offs, threads=8)
    assert out.read_bytes() == oracle.format_ovl(ref.records)


def test_kmer_cover_probe_mode(oracle, crp177):
    """--test-kmer-cover (src/Project4.scala:299-320): unique k-mers and the bucket-size histogram for k = 0..25."""
    bases, offs = crp177
    lens = np.diff(offs.astype(np.int64))
    want = []
    for k in range(26):
        if k == 0:
            uniq, hist = 1, {int((lens + 1).sum()): 1}
        else:
            r = oracle.run(oracle.default_settings(k=k), bases, offs, threads=4, keep_stages=True)
            _, sizes = np.unique(r.kmer_code, return_counts=True)
            uniq = len(sizes)
            vals, cnt = np.unique(sizes, return_counts=True)
            hist = dict(zip(vals.tolist(), cnt.tolist()))
            assert uniq == r.stats["n_buckets"]
        ratio = np.float32(uniq) / np.float32(4.0 ** k)
        want += ["Kmer Size : %d" % k, "  uniques : %d" % uniq, "  ratio   : %s" % _java_float(ratio),
                 "  [ number of collisions -> count of seqs with that many collisions ] :"]
        want += ["          [%d -> %d]" % (s, hist[s]) for s in sorted(hist)] + [""]
    p = cli("-i", os.path.join(GOLDEN, "crp177.seq"), "--test-kmer-cover")
    assert p.returncode == 0, p.stderr
    assert p.stdout.decode().split("\n")[:-1] == want


def test_bench_align_mode(oracle, crp177):
    """--bench-align (src/Project4.scala:469-481): eight lines, quadratic x 4 then linear x 4, every dispatched pair."""
    bases, offs = crp177
    n = oracle.run(oracle.default_settings(filter_valid=0), bases, offs, threads=4).stats["n_dispatched"]
    p = cli("-i", os.path.join(GOLDEN, "crp177.seq"), "--bench-align")
    assert p.returncode == 0, p.stderr
    lines = [l for l in p.stdout.decode().split("\n") if l]
    names = ["%s threaded %s %s" % (t, a, b) for a in ("quad", "linear") for t in ("single", "multi") for b in ("single", "block")]
    assert len(lines) == 8
    for line, name in zip(lines, names):
        assert line.startswith("Calculated %d %s alignments in " % (n, name)) and line.endswith(" milliseconds."), line
    # --bench-align-quick: debugStop = 500 and guards that never fire (src/Project4.scala:461-465, :611, :642): zero alignments
    p = cli("-i", os.path.join(GOLDEN, "crp177.seq"), "--bench-align-quick")
    assert p.returncode == 0, p.stderr
    lines = [l for l in p.stdout.decode().split("\n") if l]
    assert len(lines) == 8
    for line, name in zip(lines, names):
        assert line.startswith("Calculated 0 %s alignments in " % name) and line.endswith(" milliseconds."), line


def test_overlaps_probe_mode(oracle, crp177):
    """--test-overlaps (src/Project4.scala:484-504): every dispatched pair in dispatch order, duds as BioLibs.dud."""
    bases, offs = crp177
    ref = oracle.run(oracle.default_settings(filter_valid=0), bases, offs, threads=4).records
    p = cli("-i", os.path.join(GOLDEN, "crp177.seq"), "--test-overlaps")
    assert p.returncode == 0, p.stderr
    assert p.stdout.decode() == want_test_overlaps(ref, bases, offs)


def test_bench_kmer_gen_mode(oracle, crp177):
    """--bench-kmer-gen (src/Project4.scala:324-349): two lines, KmerData.size unique k-mers from SequenceData.size reads."""
    import re
    bases, offs = crp177
    path = os.path.join(GOLDEN, "crp177.seq")
    for flags, k in (([], 12), (["-k", "16"], 16)):
        uniq = oracle.run(oracle.default_settings(k=k), bases, offs, threads=4).stats["n_buckets"]
        p = cli("-i", path, "--bench-kmer-gen", *flags)
        assert p.returncode == 0, p.stderr
        head = "Generated %d unique kmers from %d sequences from %s " % (uniq, len(offs) - 1, path)
        m = re.fullmatch(r"\n" + re.escape(head) + r"sequentially in \d+ milliseconds\.\n\n"
                         + re.escape(head) + r"in parellel in \d+ milliseconds\.\n\n", p.stdout.decode())
        assert m, p.stdout

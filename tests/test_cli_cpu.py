"""hovl_cli (the Project4 command line over the C ABI) and the native host I/O of libhovl.so, on the CPU:
flag parsing (src/Project4.scala:100-259), readSeq (src/BioLibs.scala:26-50), readHOXD (:66-114) and
the calcOverlaps file writer (src/Project4.scala:795-825).  Nothing here computes on a device."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT

GOLD = os.path.join(ROOT, "tests", "golden")


def cli(*args):
    import hovl
    assert os.path.exists(hovl.CLI_PATH), "hovl_cli not built: run __graft_entry__.build()"
    return subprocess.run([hovl.CLI_PATH] + list(args), capture_output=True, timeout=120)


def settings_of(*flags):
    p = cli("-i", "reads.fa", "--hovl-print-settings", *flags)
    assert p.returncode == 0, p.stderr
    fields = dict(kv.split("=", 1) for kv in p.stdout.decode().split())
    return fields, bytes.fromhex(fields["settings"])


def test_defaults_cross_the_abi_like_the_mirror():
    import hovl
    f, raw = settings_of()
    d = hovl.default_settings()
    d.filter_valid = 1
    assert raw == bytes(d)
    assert f["action"] == "calc-overlaps" and f["output"] == ""


def test_flags_mirror_readargs(oracle):
    import hovl
    _, raw = settings_of("-k", "16", "--min-identity", "95", "--max-ignore", "-60", "-gO", "150", "--gap-extend", "-7",
                         "--kmer-edge", "0.3", "--kmer-center", "-0.25", "--min-overlap", "-33", "--min-collisions", "3",
                         "--max-collisions", "99", "--match", "5", "--mismatch", "4", "--st-hash", "--single-align",
                         "-m", os.path.join(GOLD, "HOXD1.txt"))
    d = oracle.default_settings(k=16, min_identity=95, max_ignore=60, gap_open=-150, gap_extend=-7, kmer_edge=0.3,
                                kmer_center=0.25, min_overlap=33, min_coll=3, max_coll=99, filter_valid=1)
    assert raw == bytes(d)
    # last flag wins; --quadratic-align selects the second aligner, --linear-align switches back
    _, raw = settings_of("-k", "10", "-k", "14", "--quadratic-align")
    s = hovl.Settings.from_buffer_copy(raw)
    assert (s.k, s.aligner) == (14, 1)
    _, raw = settings_of("--quadratic-align", "--linear-align")
    assert hovl.Settings.from_buffer_copy(raw).aligner == 0
    # 0.98 given as a fraction stays; a value >= 1 is a percentage, scaled in binary32
    _, raw = settings_of("--min-identity", "0.97")
    assert hovl.Settings.from_buffer_copy(raw).min_identity_f32 == hovl.f32_bits(0.97)
    _, raw = settings_of("--min-identity", "1")
    assert hovl.Settings.from_buffer_copy(raw).min_identity_f32 == int((np.float32(1) * np.float32(0.01)).view(np.uint32))


def test_modes_and_errors():
    f, _ = settings_of("--test-block-dispatch", "--test-overlaps", "-o", "out.ovl")
    assert f["action"] == "test-overlaps" and f["output"] == "out.ovl"
    p = cli("-i", "x", "--no-such-flag")
    assert p.returncode == 1 and p.stdout == b"Invalid Argument : --no-such-flag\nExiting Program.\n"
    p = cli("-k", "12")
    assert p.returncode == 255 and p.stdout == b"No input file specified\n"
    p = cli("-k", "twelve", "-i", "x")
    assert p.returncode == 1 and b"NumberFormatException" in p.stderr
    p = cli("-i", "x", "-k")
    assert p.returncode == 1 and b"ArrayIndexOutOfBoundsException" in p.stderr
    p = cli("--help", "--no-such-flag")
    assert p.returncode == 0 and b"Project 4 : Sequence Overlapper" in p.stdout
    p = cli("-i", os.path.join(GOLD, "crp177.seq"), "--test-alignment")
    assert p.returncode == 2 and b"not served" in p.stderr
    # extension flags: one handle over several GPUs switches to the wide pair key and canonical record order
    f, raw = settings_of("--hovl-devices", "0,1,3")
    assert f["devices"] == "0,1,3"
    import hovl
    s = hovl.Settings.from_buffer_copy(raw)
    assert (s.order, s.id_mode) == (1, 1)
    f, raw = settings_of("--hovl-device", "2")
    s = hovl.Settings.from_buffer_copy(raw)
    assert f["devices"] == "2" and (s.order, s.id_mode) == (0, 0)
    p = cli("-i", "x", "--hovl-devices", "0,one")
    assert p.returncode == 1 and b"NumberFormatException" in p.stderr


def test_native_readseq_equals_mirror_and_oracle(oracle, tmp_path):
    import hovl
    from hovl import io as hio
    for name in ("crp177.seq", "c_ruddii.fasta"):
        path = os.path.join(GOLD, name)
        b0, o0 = hio.read_seq(path)
        b1, o1 = oracle.read_fasta(path)
        with hovl.Reads(path) as r:
            assert r.n_reads == len(o0) - 1
            assert np.array_equal(r.bases, b0) and np.array_equal(r.offsets, o0)
            assert np.array_equal(r.bases, b1) and np.array_equal(r.offsets, o1)
    # lower case, CRLF, a bare CR, an empty record, no newline at EOF, header text ignored
    p = tmp_path / "odd.fa"
    p.write_bytes(b">one two\r\nacgt\r\nNNac\r>\n>three\nTT\rgg\n>last\nA")
    with hovl.Reads(str(p)) as r:
        assert bytes(r.bases) == b"ACGTNNACTTGGA"
        assert list(r.offsets) == [0, 8, 8, 12, 13]
    p.write_bytes(b"acgt\n>x\nAC\n")
    with pytest.raises(hovl.HovlError) as e:
        hovl.Reads(str(p))
    assert "Invalid Sequence File" in str(e.value)
    p.write_bytes(b"")
    with pytest.raises(hovl.HovlError):
        hovl.Reads(str(p))
    p.write_bytes(b">only header")
    with hovl.Reads(str(p)) as r:
        assert r.n_reads == 1 and list(r.offsets) == [0, 0]
    with pytest.raises(hovl.HovlError):
        hovl.Reads(str(tmp_path / "missing.fa"))


def test_native_readhoxd(tmp_path):
    import hovl
    from hovl import io as hio
    m = hovl.read_hoxd_native(os.path.join(GOLD, "HOXD1.txt"))
    assert m == hio.read_hoxd(os.path.join(GOLD, "HOXD1.txt"))
    assert m[0] == [91, -114, -31, -123] and m[3][3] == 91
    d = hovl.default_settings()
    assert m == [[d.cost[i][j] for j in range(4)] for i in range(4)]      # HOXD1 is the built-in matrix
    with pytest.raises(hovl.HovlError):                                   # the pair-list form: Integer.parseInt("C=-114")
        hovl.read_hoxd_native(os.path.join(GOLD, "HOXD2.txt"))
    p = tmp_path / "m.txt"
    p.write_bytes(b"#t\n-,t,g,c,a\nc,1,2,3,4\n\nA,9,9,9,9\n")            # permuted, lower case; rows after the empty line ignored
    m = hovl.read_hoxd_native(str(p))
    assert m[1] == [4, 3, 2, 1] and m[0] == [0, 0, 0, 0]
    p.write_bytes(b"#t\n-,A,C,G,T\nA, 91,1,1,1\n")                        # parseInt does not trim
    with pytest.raises(hovl.HovlError):
        hovl.read_hoxd_native(str(p))


def test_parse_int_is_integer_parseint():
    import hovl
    L = hovl.lib()
    v = C.c_int32()
    for text, want in (("12", 12), ("-7", -7), ("+5", 5), ("2147483647", 2147483647), ("-2147483648", -2147483648), ("007", 7)):
        assert L.hovl_parse_int(text.encode(), C.byref(v)) == 0 and v.value == want, text
    for text in ("", "-", "+", " 1", "1 ", "1.0", "2147483648", "-2147483649", "0x10", "1e3"):
        assert L.hovl_parse_int(text.encode(), C.byref(v)) != 0, text


def test_write_ovl_replaces_file(oracle, tmp_path):
    import hovl
    z = np.load(os.path.join(GOLD, "crp177_oracle_k12.npz"))
    recs = np.ascontiguousarray(z["records"]).view(hovl.RECORD_DTYPE).reshape(-1)
    want = open(os.path.join(GOLD, "crp177_oracle_k12.ovl"), "rb").read()
    out = tmp_path / "x.ovl"
    out.write_bytes(b"stale contents that are longer than nothing")
    assert hovl.lib().hovl_write_ovl(str(out).encode(), recs.ctypes.data, len(recs)) == 0
    assert out.read_bytes() == want


def test_fasta_modes(tmp_path):
    p = cli("-i", os.path.join(GOLD, "crp177.seq"), "--test-fasta-read")
    assert p.returncode == 0
    lines = p.stdout.decode().split("\n")
    assert lines[0] == "" and lines[1] == "id : 1" and lines[2].startswith("seq: ") and lines[3] == ""
    assert sum(1 for ln in lines if ln.startswith("id : ")) == 10
    p = cli("-i", os.path.join(GOLD, "crp177.seq"), "--bench-fasta-read")
    assert p.returncode == 0 and p.stdout.startswith(b" Read 223 sequences from ")
    p = cli("-i", str(tmp_path / "nope.fa"), "--bench-fasta-read")
    assert p.returncode == 1


def test_every_host_reads_malformed_files_the_same_way(tmp_path):
    """hovl.io (the Python host), hovl.Reads / read_hoxd_native and hovl_cli share ONE reader (csrc/hostio.cpp), so a
    file that one host accepts no other rejects: numbers are not trimmed (Integer.parseInt(row(i)),
    src/BioLibs.scala:91), a row longer than the header is an error (ArrayIndexOutOfBounds, :90), a lone '\\r' ends a
    line (BufferedReader.readLine)."""
    import hovl
    from hovl import io as hio
    fa = tmp_path / "cr.fa"
    fa.write_bytes(b">a\racgt\rTT\r>b\rgg")                              # classic-Mac line ends only
    b, o = hio.read_seq(str(fa))
    assert bytes(b) == b"ACGTTTGG" and list(o) == [0, 6, 8]
    p = cli("-i", str(fa), "--test-fasta-read")
    assert p.returncode == 0 and b"id : 1\nseq: ACGTTT\n" in p.stdout and b"id : 2\nseq: GG\n" in p.stdout
    b[0] = ord("N")                                                      # owned arrays, not views of the reader's buffer
    fa.write_bytes(b"acgt\n")
    with pytest.raises(ValueError) as e:
        hio.read_seq(str(fa))
    assert "Invalid Sequence File" in str(e.value)
    assert cli("-i", str(fa), "--test-fasta-read").returncode != 0
    reads = tmp_path / "r.fa"
    reads.write_bytes(b">1\nACGTACGTACGTACGTACGT\n>2\nCGTACGTACGTACGTACGTA\n")
    bad = {"untrimmed number": b"#t\n-,A,C,G,T\nA, 91,-114,-31,-123\n",
           "row longer than its header": b"#t\n-,A,C,G\nA,91,-114,-31,-123\n",
           "pair-list form": b"#t\nA,A=91\nA,C=-114\n",
           "unknown base label": b"#t\n-,A,C,G,T\nN,1,2,3,4\n"}
    for what, text in bad.items():
        m = tmp_path / "m.txt"
        m.write_bytes(text)
        with pytest.raises(ValueError):
            hio.read_hoxd(str(m))
        with pytest.raises(hovl.HovlError):
            hovl.read_hoxd_native(str(m))
        p = cli("-i", str(reads), "-m", str(m), "--test-fasta-read")
        assert p.returncode != 0, what
    m = tmp_path / "ok.txt"
    m.write_bytes(b"#t\r\n - , a ,C,G,T\r\n a ,5,-4,-4,-4\r\nC,-4,5,-4,-4\r\n")   # labels ARE trimmed, CRLF ends lines
    assert hio.read_hoxd(str(m))[0] == [5, -4, -4, -4] and hio.read_hoxd(str(m)) == hovl.read_hoxd_native(str(m))


def test_test_overlaps_text_and_java_floats(oracle):
    """hovl_format_test_overlaps (what hovl_cli --test-overlaps prints) against a line-by-line restatement of testOverlaps
    (src/Project4.scala:484-504) on the oracle's unfiltered records of crp177: duds, invalid and valid alignments, both
    signs of ahg.  hovl_format_float against java.lang.Float.toString."""
    import hovl
    from conftest import java_float, want_test_overlaps
    from hovl import io as hio
    bases, offs = hio.read_seq(os.path.join(GOLD, "crp177.seq"))
    recs = oracle.run(oracle.default_settings(filter_valid=0), bases, offs, threads=4).records
    assert (recs["flags"] & 1).any() and (recs["ahg"] > 0).any() and not (recs["flags"] & 4).all()
    got = hovl.format_test_overlaps(recs, bases, offs).decode()
    assert got == want_test_overlaps(recs, bases, offs)
    assert got.count(" Overlap ") == len(recs) and "   Valid?  : true\n" in got and " : 0 <-> 0\n" in got
    quad = oracle.run(oracle.default_settings(filter_valid=0, aligner=1), bases, offs, threads=4).records
    assert (quad["ahg"] < 0).any() and (quad["bhg"] < 0).any()           # --quadratic-align: the other branch of the printout
    assert hovl.format_test_overlaps(quad, bases, offs).decode() == want_test_overlaps(quad, bases, offs)
    assert hovl.format_test_overlaps(recs[:0], bases, offs) == b""
    known = {1.0: "1.0", 0.0: "0.0", 0.5: "0.5", 0.1: "0.1", 100.0: "100.0", 1e-4: "1.0E-4", 1e7: "1.0E7", 0.001: "0.001",
             9999999.0: "9999999.0", 123456.7: "123456.7", 2.0 / 3.0: "0.6666667", 0.98039216: "0.98039216",
             3.4028235e38: "3.4028235E38", -2.5: "-2.5"}
    for x, text in known.items():
        assert hovl.format_float(x) == text, x
    assert hovl.format_float(float("nan")) == "NaN" and hovl.format_float(float("inf")) == "Infinity"
    rng = np.random.default_rng(5)
    for c, e in zip(rng.integers(0, 2000, 300), rng.integers(0, 60, 300)):
        if c + e:
            x = np.float32(c) / (np.float32(c) + np.float32(e))
            assert hovl.format_float(x) == java_float(x), (c, e)

"""The reference-side binding a maintainer adds, kept as FILES so that something can check
them: jni/hovl_jni.c is type-checked against a stub jni.h (no JDK in the image), host/*.scala is checked against the
reference's signatures it calls (no scalac in the image), and scripts/pipeline_project.sh — `rake pipeline:project`
(Rakefile.rb:156-215) with hovl_cli in the overlapper slot — runs end to end through the reference's AMOS tools where
those exist (this container), on the committed .ovl that the GPU tests prove equal to hovl_run's output."""
import os
import re
import subprocess

import pytest

from conftest import GOLDEN, ROOT

AMOS = "/root/reference/amos"
TOOLS = ["toAmos_new", "bank-transact", "tigger", "make-consensus", "bank2fasta"]
HAVE_AMOS = all(os.access(os.path.join(AMOS, t), os.X_OK) for t in TOOLS)


def test_jni_shim_type_checks_against_the_c_abi():
    cmd = ["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-Wno-comment", "-fsyntax-only",
           "-I" + os.path.join(ROOT, "tests", "cpu", "jni_stub"), "-I" + os.path.join(ROOT, "include"),
           os.path.join(ROOT, "jni", "hovl_jni.c")]
    p = subprocess.run(cmd, capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    src = open(os.path.join(ROOT, "jni", "hovl_jni.c")).read()
    native = set(re.findall(r"Java_HovlNative_(\w+)", src))
    scala = set(re.findall(r"@native def (\w+)", open(os.path.join(ROOT, "host", "HovlNative.scala")).read()))
    assert native == scala == {"create", "createMulti", "destroy", "run", "formatOvl"}


def test_settings_packing_matches_the_struct_layout():
    """NativeOverlap.pack writes hovl_settings by int index; the indices must be the struct's field offsets / 4."""
    import ctypes as C
    import hovl
    src = open(os.path.join(ROOT, "host", "NativeOverlap.scala")).read()
    idx = {name: getattr(hovl.Settings, name).offset // 4 for name, _ in hovl.Settings._fields_}
    assert C.sizeof(hovl.Settings) == 32 * 4 and "new Array[Int](32)" in src
    want = {"k": "s.kmerSize", "gap_open": "s.gapOpen", "gap_extend": "s.gapExtend", "min_overlap": "s.minOverlap",
            "min_identity_f32": "floatToRawIntBits(s.minIdentity)", "max_ignore_f32": "floatToRawIntBits(s.maxIgnore)",
            "min_coll": "s.minCollisions", "max_coll": "s.maxCollisions", "head_f32": "floatToRawIntBits(s.kmerHeadEdge)",
            "tail_f32": "floatToRawIntBits(s.kmerTailEdge)", "mid_lead_f32": "floatToRawIntBits(s.kmerMidLeadEdge)",
            "mid_tail_f32": "floatToRawIntBits(s.kmerMidTailEdge)"}
    for field, expr in want.items():
        assert re.search(r"a\(%d\)\s*=\s*%s" % (idx[field], re.escape(expr)), src), (field, idx[field])
    assert idx["cost"] == 1 and "a(1 + 4 * i + j) = s.costFunc(Bases(i), Bases(j))" in src
    for field in ("aligner", "order", "id_mode", "filter_valid"):
        assert re.search(r"a\(%d\) = " % idx[field], src), field
    assert (idx["aligner"], idx["order"], idx["id_mode"], idx["filter_valid"]) == (28, 29, 30, 31)


@pytest.mark.skipif(not os.path.isdir("/root/reference/src"), reason="the reference checkout is not on this machine")
def test_scala_host_calls_the_reference_with_its_real_signatures():
    obj = open("/root/reference/src/ObjectStore.scala").read()
    prj = open("/root/reference/src/Project4.scala").read()
    host = open(os.path.join(ROOT, "host", "NativeOverlap.scala")).read() + open(os.path.join(ROOT, "host", "Project4Native.scala")).read()
    # Alignment(sA, sB, oA: String, oB: String, st, en, cor, err): 8 arguments, two of them strings
    sig = re.search(r"class Alignment\((.*?)\)\s*\{", obj, re.S).group(1)
    assert len(sig.split(",")) == 10 and "oA : String" in sig          # 8 parameters; the two tuples add a comma each
    call = re.search(r"new Alignment\((.*?)\n", host).group(1)
    assert call.count("pad") == 2                                        # placeholder alignA / alignB of length c + e
    assert "alignA.length >= settings.minOverlap" in obj
    # calcOverlaps takes (alignments, settings) and writes to Project4.output
    assert re.search(r"def calcOverlaps\(alignments : Seq\[Alignment\],settings : AlignSettings\)", prj)
    assert re.search(r"Project4\.calcOverlaps\(NativeOverlap\.align\([^)]*\), settings\)", host)
    for name in ("readArgs", "input", "output", "action", "fdAlign", "debug"):
        assert re.search(r"(def|var) %s\b" % name, prj), name


@pytest.mark.skipif(not HAVE_AMOS, reason="the reference's AMOS binaries are not on this machine")
@pytest.mark.parametrize("fixture,flags", [("crp177_oracle_k12.ovl", []), ("crp177_oracle_k16.ovl", ["-k", "16"])])
def test_pipeline_script_assembles_the_golden_contig(tmp_path, fixture, flags):
    env = dict(os.environ, HOVL_OVL=os.path.join(GOLDEN, fixture))
    work = str(tmp_path / "work")
    p = subprocess.run([os.path.join(ROOT, "scripts", "pipeline_project.sh"), os.path.join(GOLDEN, "crp177.seq"), AMOS, work,
                        "--"] + flags, capture_output=True, text=True, env=env, timeout=900)
    assert p.returncode == 0, p.stderr[-2000:]
    assert "Overlap Time" in p.stdout and "contigs: 1" in p.stdout
    got = open(os.path.join(work, "crp177.fasta"), "rb").read()
    assert got == open(os.path.join(GOLDEN, "crp177.fasta"), "rb").read()   # amos/small/crp177.fasta: 1 contig, 1,209 bp

"""Downstream functional pin: the .ovl this path emits, fed through the reference's own AMOS
tools exactly as `rake pipeline:project` does (Rakefile.rb:164-198: toAmos_new -> bank-transact -> tigger ->
make-consensus -e 0.04 -o 40 -B -> bank2fasta), must assemble to the reference's golden contig.

The AMOS tools are prebuilt ELF binaries under /root/reference/amos (they are not part of this repo and are not
copied into it): the test runs where they exist (this container) and is skipped elsewhere (the GPU box).  The
.ovl bytes are the committed oracle fixtures, which the GPU parity tests compare hovl_run / hovl_cli against."""
import os
import shutil
import subprocess

import pytest

from conftest import GOLDEN

AMOS = "/root/reference/amos"
TOOLS = ["toAmos_new", "bank-transact", "tigger", "make-consensus", "bank2fasta"]

pytestmark = pytest.mark.skipif(not all(os.access(os.path.join(AMOS, t), os.X_OK) for t in TOOLS),
                                reason="the reference's AMOS binaries are not on this machine")


def _sh(cmd, cwd, stdout=None):
    p = subprocess.run(cmd, cwd=cwd, stdout=stdout or subprocess.PIPE, stderr=subprocess.PIPE, timeout=900)
    assert p.returncode == 0, (cmd, p.stderr[-2000:])


def _assemble(tmp, seq_path, ovl_bytes):
    seq = os.path.join(tmp, "reads.seq")
    shutil.copy(seq_path, seq)
    with open(os.path.join(tmp, "reads.ovl"), "wb") as f:
        f.write(ovl_bytes)
    _sh([os.path.join(AMOS, "toAmos_new"), "-s", "reads.seq", "-b", "reads.bnk"], tmp)
    _sh([os.path.join(AMOS, "bank-transact"), "-b", "reads.bnk", "-m", "reads.ovl"], tmp)
    _sh([os.path.join(AMOS, "tigger"), "-b", "reads.bnk"], tmp)
    _sh([os.path.join(AMOS, "make-consensus"), "-e", "0.04", "-o", "40", "-B", "-b", "reads.bnk"], tmp)
    with open(os.path.join(tmp, "out.fasta"), "wb") as f:
        _sh([os.path.join(AMOS, "bank2fasta"), "-b", "reads.bnk"], tmp, stdout=f)
    return open(os.path.join(tmp, "out.fasta"), "rb").read()


@pytest.mark.parametrize("fixture", ["crp177_oracle_k12.ovl", "crp177_oracle_k16.ovl"])
def test_crp177_overlaps_assemble_to_the_golden_contig(tmp_path, fixture):
    ovl = open(os.path.join(GOLDEN, fixture), "rb").read()
    got = _assemble(str(tmp_path), os.path.join(GOLDEN, "crp177.seq"), ovl)
    want = open(os.path.join(GOLDEN, "crp177.fasta"), "rb").read()       # amos/small/crp177.fasta: 1 contig, 1,209 bp
    assert got == want


def test_quadratic_aligner_overlaps_assemble_too(tmp_path, oracle, crp177):
    bases, offs = crp177
    r = oracle.run(oracle.default_settings(aligner=1), bases, offs, threads=4)
    got = _assemble(str(tmp_path), os.path.join(GOLDEN, "crp177.seq"), oracle.format_ovl(r.records))
    assert got == open(os.path.join(GOLDEN, "crp177.fasta"), "rb").read()


def test_record_order_does_not_matter_downstream(tmp_path, oracle, crp177):
    """tigger does not depend on the order of the OVL records: canonical (a, b) order assembles the same contig,
    which is why order = 1 (what reads beyond 32767 ids and the sharded path return) is a drop-in too."""
    bases, offs = crp177
    r = oracle.run(oracle.default_settings(order=1, id_mode=1), bases, offs, threads=4)
    got = _assemble(str(tmp_path), os.path.join(GOLDEN, "crp177.seq"), oracle.format_ovl(r.records))
    assert got == open(os.path.join(GOLDEN, "crp177.fasta"), "rb").read()

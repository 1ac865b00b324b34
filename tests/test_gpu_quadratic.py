"""--quadratic-align on a GPU (quad_kernel, csrc/quad.cu) through the C ABI against the oracle's literal
restatement of generateLocalAlignmentSet (src/BioLibs.scala:267-368): all 13 record fields, .ovl bytes."""
import os
import subprocess

import numpy as np
import pytest

from conftest import GOLDEN, synth_reads

pytestmark = pytest.mark.gpu

FIELDS = ["a", "b", "flags", "start_i", "start_j", "end_i", "end_j", "correct", "error", "score", "ahg", "bhg"]


def _settings(oracle, **kw):
    import hovl
    a, b = hovl.default_settings(aligner=1, **kw), oracle.default_settings(aligner=1, **kw)
    assert bytes(a) == bytes(b)
    return a, b


@pytest.mark.parametrize("k", [12, 16])
def test_stage_align_all_candidate_pairs(hovl_ctx, oracle, crp177, k):
    bases, offs = crp177
    hs, os_ = _settings(oracle, k=k, min_coll=1, max_coll=2**31 - 1, filter_valid=0)
    r = oracle.run(os_, bases, offs, threads=4, keep_stages=True)
    ref, _ = oracle.align_pairs(os_, bases, offs, r.disp_a, r.disp_b, threads=8)
    got = hovl_ctx.stage_align(hs, bases, offs, r.disp_a, r.disp_b)
    for f in FIELDS:
        assert np.array_equal(got[f], ref[f]), f


@pytest.mark.parametrize("L,indel,ragged,go,ge,cm", [(300, 0.01, 150, -200, -20, None), (1000, 0.004, 300, -200, -20, None),
                                                      (130, 0.02, 100, 0, 0, (1, -1)), (700, 0.01, 650, -3, 0, (3, -2)),
                                                      (2100, 0.002, 50, -8, -2, (5, -4))])
def test_stage_align_column_tiles_ties_ragged(hovl_ctx, oracle, L, indel, ragged, go, ge, cm):
    """Several 128-column tiles, reads of very different lengths in one call, tie-heavy cost settings."""
    n = 24 if L <= 1000 else 10
    bases, offs = synth_reads(L + 3, n, L, 2 * L, sub=0.02, indel=indel, ragged=ragged)
    hs, os_ = _settings(oracle, filter_valid=0, min_identity=0.9, max_ignore=100000)
    for s in (hs, os_):
        s.gap_open, s.gap_extend = go, ge
        if cm:
            for i in range(4):
                for j in range(4):
                    s.cost[i][j] = cm[0] if i == j else cm[1]
    pa = np.repeat(np.arange(1, n + 1), n).astype(np.uint32)
    pb = np.tile(np.arange(1, n + 1), n).astype(np.uint32)
    ref, cells = oracle.align_pairs(os_, bases, offs, pa, pb, threads=8)
    got = hovl_ctx.stage_align(hs, bases, offs, pa, pb)
    for f in FIELDS:
        assert np.array_equal(got[f], ref[f]), f
    assert (ref["flags"] & 2).sum() > 0


@pytest.mark.parametrize("kw", [{}, {"k": 16}, {"order": 1, "id_mode": 1, "filter_valid": 0}])
def test_run_crp177_byte_identical_ovl(hovl_ctx, oracle, crp177, kw):
    bases, offs = crp177
    hs, os_ = _settings(oracle, **kw)
    res = hovl_ctx.run(hs, bases, offs)
    ref = oracle.run(os_, bases, offs, threads=8)
    assert len(res.records) == len(ref.records) > 2000
    for f in FIELDS + ["count"]:
        assert np.array_equal(res.records[f], ref.records[f]), f
    assert res.ovl() == oracle.format_ovl(ref.records)
    assert res.stats["dp_cells"] == ref.stats["dp_cells"] == 100 * 100 * res.stats["n_dispatched"]


def test_run_synthetic_1kb_and_cli(hovl_ctx, oracle, tmp_path):
    import hovl
    from hovl import io as hio
    bases, offs = synth_reads(77, 300, 1000, 12000, sub=0.01)
    hs, os_ = _settings(oracle, k=16, max_ignore=1000, max_coll=2**31 - 1)
    res = hovl_ctx.run(hs, bases, offs)
    ref = oracle.run(os_, bases, offs, threads=8)
    assert len(ref.records) > 500
    for f in FIELDS + ["count"]:
        assert np.array_equal(res.records[f], ref.records[f]), f
    fa, out = tmp_path / "r.fa", tmp_path / "r.ovl"
    hio.write_fasta(str(fa), bases, offs)
    p = subprocess.run([hovl.CLI_PATH, "-i", str(fa), "-o", str(out), "-k", "16", "--max-ignore", "1000",
                        "--max-collisions", "2147483647", "--quadratic-align"], capture_output=True, timeout=300)
    assert p.returncode == 0, p.stderr
    assert out.read_bytes() == oracle.format_ovl(ref.records)


def test_no_positive_cell_bad_base_empty(hovl_ctx, oracle):
    bases = np.frombuffer(b"AAAAAAAAAAAAAAAAAAAA" + b"CCCCCCCCCCCCCCCCCCCCCCCCC" + b"ACGTNACGTACGTACGTACGT", np.uint8)
    offs = np.array([0, 20, 45, 66], np.uint64)
    pa = np.array([1, 2, 1, 3, 2, 3], np.uint32)
    pb = np.array([2, 1, 3, 1, 3, 2], np.uint32)
    hs, os_ = _settings(oracle, filter_valid=0)
    ref, _ = oracle.align_pairs(os_, bases, offs, pa, pb, threads=1)
    got = hovl_ctx.stage_align(hs, bases, offs, pa, pb)
    for f in FIELDS:
        assert np.array_equal(got[f], ref[f]), f

"""--quadratic-align (generateLocalAlignment, src/BioLibs.scala:171-368) on the CPU: the oracle's literal
full-matrix restatement against the kernel's own lane math (csrc/quad_core.cuh run by the host lane
emulator tests/cpu/quad_emul.cpp: column tiles, skewed wavefront, forward-carried trace tags)."""
import numpy as np
import pytest

from conftest import synth_reads


def _all_pairs(n):
    a, b = np.meshgrid(np.arange(1, n + 1), np.arange(1, n + 1), indexing="ij")
    keep = a != b
    return a[keep].astype(np.uint32), b[keep].astype(np.uint32)


def _same(ref, got, tag):
    for f in ref.dtype.names:
        if f != "count":
            assert np.array_equal(ref[f], got[f]), (f, tag)


def test_quadratic_oracle_on_crp177_is_a_subset_of_the_amos_golden(oracle, crp177):
    """The only reference-side pin the path has: every (rds, ahg, bhg) emitted on
    crp177.seq is in AMOS's own crp177.ovl, for the quadratic aligner as for the fast one."""
    import os
    import re
    from conftest import GOLDEN
    bases, offs = crp177
    r = oracle.run(oracle.default_settings(aligner=1), bases, offs, threads=4)
    txt = open(os.path.join(GOLDEN, "crp177.ovl")).read()
    gold = {tuple(int(x) for x in m) for m in re.findall(r"rds:(\d+),(\d+)\nscr:\d+\nahg:(-?\d+)\nbhg:(-?\d+)", txt)}
    mine = {(int(x["a"]), int(x["b"]), int(x["ahg"]), int(x["bhg"])) for x in r.records}
    assert len(mine) > 2000 and mine <= gold


@pytest.mark.parametrize("k", [12, 16])
def test_lane_math_on_crp177_all_candidate_pairs(oracle, quad_emul, crp177, k):
    bases, offs = crp177
    s = oracle.default_settings(k=k, min_coll=1, max_coll=2**31 - 1, filter_valid=0, aligner=1)
    r = oracle.run(s, bases, offs, threads=4, keep_stages=True)
    ref, cells = oracle.align_pairs(s, bases, offs, r.disp_a, r.disp_b, threads=4)
    for D in (1, 4):
        got, cells2 = quad_emul(s, bases, offs, r.disp_a, r.disp_b, D)
        _same(ref, got, D)
        assert cells == cells2
    assert (ref["flags"] & 2).sum() > 1000


@pytest.mark.parametrize("L,indel,ragged", [(300, 0.01, 150), (130, 0.02, 100), (520, 0.004, 30)])
def test_lane_math_across_column_tiles(oracle, quad_emul, L, indel, ragged):
    """Reads longer than one tile (32*D columns): the boundary column between tiles, ragged lengths, indels."""
    bases, offs = synth_reads(L, 14, L, 2 * L, sub=0.02, indel=indel, ragged=ragged)
    pa, pb = _all_pairs(14)
    s = oracle.default_settings(filter_valid=0, min_identity=0.9, max_ignore=100000, aligner=1)
    ref, cells = oracle.align_pairs(s, bases, offs, pa, pb, threads=4)
    for D in (1, 2, 4):
        got, cells2 = quad_emul(s, bases, offs, pa, pb, D)
        _same(ref, got, D)
        assert cells == cells2
    assert (ref["flags"] & 2).sum() > 0


def test_lane_math_ties_and_odd_costs(oracle, quad_emul):
    """Flat cost matrices and zero gap penalties make M/X/Y ties and equal maxima on several rows and
    tiles common: priority M, X, Y in the trace and strict '>' in row-major order in the arg-max."""
    bases, offs = synth_reads(11, 12, 200, 400, sub=0.08, indel=0.04, ragged=90)
    pa, pb = _all_pairs(12)
    for go, ge, cm in [(-1, -1, (2, -1)), (0, 0, (1, -1)), (-3, 0, (3, -2)), (0, -1, (1, -2)), (-200, -20, None)]:
        s = oracle.default_settings(filter_valid=0, min_identity=0.9, max_ignore=100000, aligner=1)
        s.gap_open, s.gap_extend = go, ge
        if cm:
            for i in range(4):
                for j in range(4):
                    s.cost[i][j] = cm[0] if i == j else cm[1]
        ref, _ = oracle.align_pairs(s, bases, offs, pa, pb, threads=4)
        for D in (1, 4):
            got, _ = quad_emul(s, bases, offs, pa, pb, D)
            _same(ref, got, (go, ge, D))


def test_lane_math_no_positive_cell_and_bad_base(oracle, quad_emul):
    """All-mismatch pairs never get a positive cell (the reference would index charAt(-1)): dud;
    a non-ACGT base is costFunc's MatchError: dud + BAD_BASE."""
    bases = np.frombuffer(b"AAAAAAAAAAAAAAAAAAAA" + b"CCCCCCCCCCCCCCCCCCCCCCCCC" + b"ACGTNACGTACGTACGTACGT", np.uint8)
    offs = np.array([0, 20, 45, 66], np.uint64)
    pa, pb = _all_pairs(3)
    s = oracle.default_settings(filter_valid=0, aligner=1)
    ref, _ = oracle.align_pairs(s, bases, offs, pa, pb, threads=1)
    got, _ = quad_emul(s, bases, offs, pa, pb, 4)
    _same(ref, got, "dud")
    assert (ref["flags"] & 1).all() and ((ref["flags"] & 16) != 0).sum() == 4

"""The wide-band aligner's per-lane code (csrc/dp_wb.cuh: warp per pair, skewed wavefront, packed score+priority
cells, 2-bit direction codes in 16-step blocks, packed arg-max keys, padded diagonals neutralised through the
gap-open constant) against the oracle's full-matrix DP + backward trace (src/BioLibs.scala:596-822), on the host:
tests/cpu/wb_emul.cpp runs the same source over 32 emulated lanes."""
import numpy as np
import pytest

from conftest import synth_reads


def _all_pairs(n):
    pa = np.repeat(np.arange(1, n + 1), n)
    pb = np.tile(np.arange(1, n + 1), n)
    keep = pa != pb
    return pa[keep].astype(np.uint32), pb[keep].astype(np.uint32)


def _check(oracle, wb_emul, s, bases, offs, pa, pb, D, want_valid=True):
    ref, cells = oracle.align_pairs(s, bases, offs, pa, pb, threads=4)
    got, cells2 = wb_emul(s, bases, offs, pa, pb, D)
    wide = got["flags"] == -1                     # band wider than 32 * D - 1: those pairs belong to the next class
    if wide.any():
        assert D < 8
        g2, c2 = wb_emul(s, bases, offs, pa[wide], pb[wide], D + 1)
        got[wide] = g2
        cells2 += c2
    assert (got["flags"] >= 0).all()
    for f in ref.dtype.names:
        if f != "count":
            bad = np.nonzero(ref[f] != got[f])[0]
            assert len(bad) == 0, (f, bad[:5], ref[bad[:5]], got[bad[:5]])
    assert cells == cells2
    if want_valid:
        assert (ref["flags"] & 2).sum() > 0 and (ref["flags"] & 1).sum() > 0


# (read length, identity) -> band width = max(k, floor(f32(L) * (1f - mi)) + 1); D = diagonals per lane
@pytest.mark.parametrize("L,mi,D,indel,ragged", [
    (1700, 0.98, 2, 0.004, 30),       # width 35: the narrowest wide class, 29 padded diagonals
    (1000, 0.95, 2, 0.01, 0),         # width 51
    (1000, 0.937, 2, 0.004, 100),     # width 63/64 boundary of D = 2 (ragged lengths straddle it)
    (1500, 0.94, 3, 0.01, 200),       # width ~ 79..91
    (1200, 0.90, 4, 0.01, 30),        # width 121
    (800, 0.80, 6, 0.02, 0),          # width 161 -> D = 6 (cap 191)
    (1000, 0.80, 7, 0.02, 40),        # width 193..201 -> D = 7
    (1200, 0.80, 8, 0.02, 0),         # width 241 -> D = 8
])
def test_wide_band_lane_math_on_synthetic_reads(oracle, wb_emul, L, mi, D, indel, ragged):
    bases, offs = synth_reads(L + D, 16, L, 3 * L, sub=0.01, indel=indel, ragged=ragged)
    s = oracle.default_settings(k=12, filter_valid=0, min_identity=mi, max_ignore=100000)
    pa, pb = _all_pairs(16)
    _check(oracle, wb_emul, s, bases, offs, pa, pb, D)


def test_wide_band_k5_shape(oracle, wb_emul):
    """K5's shape: 10 kb reads, 1 % substitutions, k = 16 -> width 200 (src/BioLibs.scala:619-620), D = 7."""
    bases, offs = synth_reads(5, 8, 10000, 16000, sub=0.01)
    s = oracle.default_settings(k=16, filter_valid=0, max_ignore=10000)
    pa, pb = _all_pairs(8)
    _check(oracle, wb_emul, s, bases, offs, pa, pb, 7)
    _check(oracle, wb_emul, s, bases, offs, pa[:12], pb[:12], 8)      # the same pairs in a class with more padding


@pytest.mark.parametrize("go,ge,cm", [(-1, -1, (2, -1)), (0, 0, (1, -1)), (-3, 0, (3, -2)), (-200, -20, None),
                                      (-100, -10, (50, -60)), (-1, -1, (7, 6))])
def test_wide_band_ties_and_odd_costs(oracle, wb_emul, go, ge, cm):
    """Small gap penalties and flat cost matrices make M/X/Y ties and equal maxima common: the priority bits must
    resolve them in the reference's M, X, Y order and the arg-max keys in its row-major order."""
    bases, offs = synth_reads(7, 14, 600, 1300, sub=0.05, indel=0.03, ragged=120)
    pa, pb = _all_pairs(14)
    s = oracle.default_settings(k=12, filter_valid=0, min_identity=0.9, max_ignore=100000)     # width 49..61
    s.gap_open, s.gap_extend = go, ge
    if cm:
        for i in range(4):
            for j in range(4):
                s.cost[i][j] = cm[0] if i == j else cm[1]
    _check(oracle, wb_emul, s, bases, offs, pa, pb, 2, want_valid=False)
    _check(oracle, wb_emul, s, bases, offs, pa[:40], pb[:40], 3, want_valid=False)


def test_wide_band_short_b_bad_base_and_repeats(oracle, wb_emul):
    """B shorter than the band (StringIndexOutOfBounds in the reference), non-ACGT bases (MatchError), B barely longer
    than the band (every phase-2 row range-checked) and a two-letter repeat genome (many equal maxima)."""
    rng = np.random.default_rng(3)
    reads = [rng.integers(0, 4, 1000), rng.integers(0, 4, 30), rng.integers(0, 4, 52), rng.integers(0, 2, 900) * 3,
             rng.integers(0, 2, 1000) * 3]
    reads[2][:40] = reads[0][960:1000]
    reads.append(reads[0][100:900].copy())
    offs = np.zeros(len(reads) + 1, np.uint64)
    offs[1:] = np.cumsum([len(r) for r in reads])
    bases = np.frombuffer(b"ACGT", np.uint8)[np.concatenate(reads)].copy()
    bases[int(offs[5]) + 7] = ord("N")
    s = oracle.default_settings(k=12, filter_valid=0, min_identity=0.95, max_ignore=100000)    # width 51 for L = 1000
    pa, pb = _all_pairs(len(reads))
    _check(oracle, wb_emul, s, bases, offs, pa, pb, 2, want_valid=False)

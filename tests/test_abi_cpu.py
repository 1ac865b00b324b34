"""CPU-side checks of the product library: it loads, exports every symbol include/hovl.h
declares, refuses to compute without a GPU (no CPU fallback), and its host-only pieces
(settings, OVL formatting, Trove order replay, DP lane math) agree with the oracle."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, synth_reads


def _declared_functions():
    txt = open(os.path.join(ROOT, "include", "hovl.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(hovl_[a-z_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    import hovl
    L = hovl.lib()
    names = _declared_functions()
    assert len(names) >= 16
    for n in names:
        assert hasattr(L, n), n
    assert sorted(hovl.EXPORTS) == names


def test_struct_layouts_match_header():
    import hovl
    assert C.sizeof(hovl.Settings) == 4 * (1 + 16 + 2 + 1 + 2 + 2 + 4 + 4)
    assert hovl.RECORD_DTYPE.itemsize == 52
    assert C.sizeof(hovl.Stats) == 8 * len(hovl.STATS_FIELDS)


def test_defaults_equal_oracle_defaults(oracle):
    import hovl
    a, b = hovl.default_settings(), oracle.default_settings()
    assert bytes(a) == bytes(b)
    a = hovl.default_settings(k=16, kmer_edge=0.3, kmer_center=0.25, min_identity=95, max_ignore=-60, gap_open=150)
    b = oracle.default_settings(k=16, kmer_edge=0.3, kmer_center=0.25, min_identity=95, max_ignore=-60, gap_open=-150)
    assert bytes(a) == bytes(b)
    # the four section thresholds of the defaults
    d = hovl.default_settings()
    assert (d.head_f32, d.mid_lead_f32, d.tail_f32, d.mid_tail_f32) == (0x3ECCCCCD, 0x3E99999A, 0x3F19999A, 0x3F333333)


def test_no_cpu_fallback():
    import torch
    import hovl
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(hovl.HovlError) as e:
        hovl.Context(0)
    assert e.value.code == hovl.E_NO_DEVICE


def test_format_ovl_equals_oracle(oracle):
    import hovl
    g = np.load(os.path.join(ROOT, "tests", "golden", "crp177_oracle_k12.npz"))
    recs = g["records"].view(hovl.RECORD_DTYPE).reshape(-1)
    assert hovl.format_ovl(recs) == open(os.path.join(ROOT, "tests", "golden", "crp177_oracle_k12.ovl"), "rb").read()
    mixed = recs.copy()
    mixed["flags"][::3] = 1                      # duds are not printed
    assert hovl.format_ovl(mixed) == oracle.format_ovl(mixed)


@pytest.mark.parametrize("n,seed", [(0, 0), (1, 1), (11, 2), (12, 3), (500, 4), (20000, 5), (300000, 6)])
def test_trove_order_replay_equals_oracle_model(oracle, n, seed):
    import hovl
    rng = np.random.default_rng(seed)
    if seed % 2:
        keys = rng.choice(np.arange(-2**31, 2**31 - 1, dtype=np.int64), size=n, replace=False).astype(np.int32) if n else np.zeros(0, np.int32)
    else:
        keys = rng.permutation(np.arange(n, dtype=np.int64) * 7 - 3 * n).astype(np.int32)
    L = oracle.lib()
    m = L.trove_new()
    isnew = C.c_int()
    for k in keys.tolist():
        L.trove_put_if_absent(m, k, C.byref(isnew))
        assert isnew.value == 1
    ref = (C.c_int32 * max(n, 1))()
    L.trove_iterate(m, ref, None)
    L.trove_free(m)
    got = keys[hovl.trove_order(keys)]
    assert got.tolist() == list(ref)[:n]


def _all_pairs(n):
    pa = np.repeat(np.arange(1, n + 1), n)
    pb = np.tile(np.arange(1, n + 1), n)
    keep = pa != pb
    return pa[keep].astype(np.uint32), pb[keep].astype(np.uint32)


@pytest.mark.parametrize("k,G,D", [(12, 8, 2), (16, 8, 3), (12, 16, 3)])
def test_dp_lane_math_on_crp177_all_candidate_pairs(oracle, emul, crp177, k, G, D):
    """Every candidate pair of crp177 (count filter off), duds included: the skewed-wavefront /
    forward-trace formulation used by the CUDA kernel must reproduce the reference's
    full-matrix DP + backward trace field by field."""
    bases, offs = crp177
    s = oracle.default_settings(k=k, min_coll=1, max_coll=2**31 - 1, filter_valid=0)
    r = oracle.run(s, bases, offs, threads=4, keep_stages=True)
    ref, cells = oracle.align_pairs(s, bases, offs, r.disp_a, r.disp_b, threads=4)
    got, cells2 = emul(s, bases, offs, r.disp_a, r.disp_b, G, D)
    for f in ref.dtype.names:
        if f != "count":
            assert np.array_equal(ref[f], got[f]), f
    assert cells == cells2


@pytest.mark.parametrize("L,G,D,mi,indel", [(400, 8, 2, 0.98, 0.004), (1000, 8, 3, 0.98, 0.004), (1000, 32, 4, 0.90, 0.01),
                                            (1500, 16, 4, 0.97, 0.0), (1200, 32, 8, 0.80, 0.02)])
def test_dp_lane_math_on_synthetic_reads_with_indels(oracle, emul, L, G, D, mi, indel):
    bases, offs = synth_reads(L, 24, L, 3 * L, sub=0.01, indel=indel, ragged=30)
    s = oracle.default_settings(k=12, filter_valid=0, min_identity=mi, max_ignore=100000)
    pa, pb = _all_pairs(24)
    ref, cells = oracle.align_pairs(s, bases, offs, pa, pb, threads=4)
    got, cells2 = emul(s, bases, offs, pa, pb, G, D)
    assert (got["flags"] >= 0).all()
    for f in ref.dtype.names:
        if f != "count":
            assert np.array_equal(ref[f], got[f]), f
    assert cells == cells2
    assert (ref["flags"] & 2).sum() > 0 and (ref["flags"] & 1).sum() > 0


@pytest.mark.parametrize("k", [12, 14, 16])
def test_dp_thread_per_pair_on_crp177_all_candidate_pairs(oracle, tp_emul, crp177, k):
    """The narrow-band kernel's per-thread code (packed score+priority cells, 2-bit direction
    codes, backward walk; csrc/dp_tp.cuh) against the reference's full-matrix DP + trace."""
    bases, offs = crp177
    s = oracle.default_settings(k=k, min_coll=1, max_coll=2**31 - 1, filter_valid=0)
    r = oracle.run(s, bases, offs, threads=4, keep_stages=True)
    ref, cells = oracle.align_pairs(s, bases, offs, r.disp_a, r.disp_b, threads=4)
    for fast in (0, 1, 2, 3):
        got, cells2 = tp_emul(s, bases, offs, r.disp_a, r.disp_b, fast)
        for f in ref.dtype.names:
            if f != "count":
                assert np.array_equal(ref[f], got[f]), (f, fast)
        assert fast >= 2 or cells == cells2


@pytest.mark.parametrize("L,k,mi,indel,ragged", [(400, 12, 0.98, 0.004, 30), (1000, 12, 0.98, 0.004, 30),
                                                 (1000, 16, 0.98, 0.0, 0), (700, 12, 0.97, 0.01, 200),
                                                 (1590, 12, 0.98, 0.004, 40), (300, 12, 0.90, 0.02, 100),
                                                 (120, 12, 0.98, 0.01, 100)])
def test_dp_thread_per_pair_on_synthetic_reads_with_indels(oracle, tp_emul, L, k, mi, indel, ragged):
    bases, offs = synth_reads(L + k, 24, L, 3 * L, sub=0.01, indel=indel, ragged=ragged)
    s = oracle.default_settings(k=k, filter_valid=0, min_identity=mi, max_ignore=100000)
    pa, pb = _all_pairs(24)
    ref, cells = oracle.align_pairs(s, bases, offs, pa, pb, threads=4)
    for fast in ((0, 1, 2, 3) if L <= 1023 else (0, 2, 3)):
        got, cells2 = tp_emul(s, bases, offs, pa, pb, fast)
        for f in ref.dtype.names:
            if f != "count":
                assert np.array_equal(ref[f], got[f]), (f, fast)
        assert fast >= 2 or cells == cells2
    assert (ref["flags"] & 2).sum() > 0 and (ref["flags"] & 1).sum() > 0


@pytest.mark.parametrize("L,k,mi,indel,ragged,cm,go,ge", [(1000, 16, 0.98, 0.004, 30, (5, -4), -8, -2),
                                                          (700, 12, 0.97, 0.01, 300, (1, -1), -2, -1),
                                                          (1590, 12, 0.98, 0.004, 400, (2, -3), -5, -2),
                                                          (130, 12, 0.95, 0.02, 100, (50, -60), -100, -10),
                                                          (1000, 12, 0.97, 0.006, 300, None, -200, -20),
                                                          (1000, 12, 0.972, 0.002, 500, None, -60, -5),
                                                          (900, 12, 0.98, 0.01, 0, (150, -180), -250, -30),
                                                          (1000, 12, 0.97, 0.01, 100, (7, -90), -3, 0)])
def test_dp_two_pairs_per_thread_both_phases(oracle, tp_emul, L, k, mi, indel, ragged, cm, go, ge):
    """csrc/dp_tp2.cuh: 16-bit half-word cells, two pairs of different lengths / dovetail lengths in
    one thread, arg-max by superblock replay; scores that fit 13 bits."""
    bases, offs = synth_reads(L + k + 1, 25, L, 3 * L, sub=0.01 if cm else 0.012, indel=indel, ragged=ragged)
    s = oracle.default_settings(k=k, filter_valid=0, min_identity=mi, max_ignore=100000)
    s.gap_open, s.gap_extend = go, ge
    if cm:
        for i in range(4):
            for j in range(4):
                s.cost[i][j] = cm[0] if i == j else cm[1]
    pa, pb = _all_pairs(25)
    ref, _ = oracle.align_pairs(s, bases, offs, pa, pb, threads=4)
    for fast in (0, 2, 3):
        got, _ = tp_emul(s, bases, offs, pa, pb, fast)
        for f in ref.dtype.names:
            if f != "count":
                assert np.array_equal(ref[f], got[f]), (f, fast)
    assert cm == (7, -90) or ((ref["flags"] & 2).sum() > 0 and (ref["flags"] & 1).sum() > 0)


def test_dp_thread_per_pair_ties_and_odd_costs(oracle, tp_emul):
    """Small gap penalties and a flat cost matrix make M/X/Y ties common: the packed priority
    bits must resolve them in the reference's M, X, Y order."""
    bases, offs = synth_reads(7, 20, 300, 700, sub=0.05, indel=0.03, ragged=60)
    pa, pb = _all_pairs(20)
    for go, ge, cm in [(-1, -1, (2, -1)), (0, 0, (1, -1)), (-3, 0, (3, -2)), (-200, -20, None)]:
        s = oracle.default_settings(k=12, filter_valid=0, min_identity=0.95, max_ignore=100000)
        s.gap_open, s.gap_extend = go, ge
        if cm:
            for i in range(4):
                for j in range(4):
                    s.cost[i][j] = cm[0] if i == j else cm[1]
        ref, cells = oracle.align_pairs(s, bases, offs, pa, pb, threads=4)
        for fast in (0, 1, 2, 3):
            got, cells2 = tp_emul(s, bases, offs, pa, pb, fast)
            for f in ref.dtype.names:
                if f != "count":
                    assert np.array_equal(ref[f], got[f]), (f, go, ge, fast)


@pytest.mark.parametrize("L,k,mi,sub,indel,ragged,go,ge,cm", [
    (100, 12, 0.98, 0.01, 0.0, 0, -200, -20, None), (1000, 16, 0.98, 0.01, 0.0, 0, -200, -20, None),
    (400, 12, 0.96, 0.02, 0.01, 150, -200, -20, None), (700, 12, 0.97, 0.03, 0.01, 300, 0, -1, (1, -1)),
    (300, 12, 0.95, 0.05, 0.03, 60, -3, -1, (3, -2)), (1500, 12, 0.98, 0.004, 0.002, 40, -5, -2, (2, -3)),
    (250, 14, 0.9, 0.0, 0.0, 0, -1, -1, (7, 6))])
def test_phase1_exact_prefix_probe_equals_phase1_dp(oracle, tp_emul, L, k, mi, sub, indel, ragged, go, ge, cm):
    """dp_tp.cuh probe_prefix: wherever B's first `width` bases occur in A exactly, the first occurrence IS
    phase 1's doveStart (arg-max + greedy trace of the DP), for any cost matrix that probe_valid accepts —
    low-complexity genomes (many occurrences, ties), indels, ragged lengths, tiny gap penalties."""
    n = 30
    rng = np.random.default_rng(L + k)
    bases, offs = synth_reads(L + k, n, L, 3 * L, sub=sub, indel=indel, ragged=ragged)
    if cm == (7, 6):                                   # a two-letter, repeat-rich genome: many exact occurrences
        bases = np.where(bases == ord("G"), ord("A"), np.where(bases == ord("T"), ord("C"), bases)).astype(np.uint8)
    s = oracle.default_settings(k=k, filter_valid=0, min_identity=mi, max_ignore=100000)
    s.gap_open, s.gap_extend = go, ge
    if cm:
        for i in range(4):
            for j in range(4):
                s.cost[i][j] = cm[0] if i == j else cm[1]
    pa, pb = _all_pairs(n)
    valid, ds_dp, ds_probe = tp_emul.probe(s, bases, offs, pa, pb)
    assert valid
    ran = ds_dp != -2
    hit = ran & (ds_probe >= 0)
    assert hit.sum() > 20
    assert np.array_equal(ds_dp[hit], ds_probe[hit])
    assert (ran & (ds_probe == -1)).sum() > 0          # and pairs the DP has to decide exist too
    del rng


def test_phase1_probe_refuses_cost_settings_outside_the_proof(oracle, tp_emul):
    bases, offs = synth_reads(3, 4, 100, 300)
    pa, pb = _all_pairs(4)
    for go, ge, cm in [(-200, 0, None), (0, 0, (1, -1)), (-2, -1, (1, 1)), (-2, -1, (0, -1)), (-2, -1, (-1, -3))]:
        s = oracle.default_settings(filter_valid=0)
        s.gap_open, s.gap_extend = go, ge
        if cm:
            for i in range(4):
                for j in range(4):
                    s.cost[i][j] = cm[0] if i == j else cm[1]
        valid, _, _ = tp_emul.probe(s, bases, offs, pa, pb)
        assert not valid

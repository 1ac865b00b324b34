"""GPU parity tests: the CUDA path, called through the C ABI (libhovl.so), against the CPU oracle
on the same inputs — bit-exact for every integer the path produces (k-mer codes, pair keys and
counts, dispatch order, all 13 record fields, .ovl bytes)."""
import hashlib
import os

import numpy as np
import pytest

from conftest import GOLDEN, synth_reads

pytestmark = pytest.mark.gpu

FIELDS = ["a", "b", "count", "flags", "start_i", "start_j", "end_i", "end_j", "correct", "error", "score", "ahg", "bhg"]


def _same_settings(oracle, **kw):
    import hovl
    a, b = hovl.default_settings(**kw), oracle.default_settings(**kw)
    assert bytes(a) == bytes(b)
    return a, b


def _assert_records_equal(got, ref):
    assert len(got) == len(ref)
    for f in FIELDS:
        assert np.array_equal(got[f], ref[f]), f


def _c_ruddii_reads():
    from hovl import io as hio
    contig, _ = hio.read_seq(os.path.join(GOLDEN, "c_ruddii.fasta"))
    lay = np.load(os.path.join(GOLDEN, "c_ruddii_layout.npy")).astype(np.int64)
    bases = contig[lay[:, None] + np.arange(100)[None, :]].reshape(-1)
    return np.ascontiguousarray(bases), (np.arange(len(lay) + 1) * 100).astype(np.uint64)


# ------------------------------------------------------------------ stages
@pytest.mark.parametrize("k", [12, 16, 5, 20])
def test_stage_table_crp177(hovl_ctx, oracle, crp177, k):
    """a1+a2: splitter tuples grouped by code in arrival order (BioLibs.scala:54-61, KmerTable.scala:41-53)."""
    bases, offs = crp177
    hs, os_ = _same_settings(oracle, k=k)
    code, rid, loc = hovl_ctx.stage_table(hs, bases, offs)
    r = oracle.run(os_, bases, offs, threads=4, keep_stages=True)
    assert np.array_equal(code, r.kmer_code)
    assert np.array_equal(rid, r.kmer_rid)
    assert np.array_equal(loc, r.kmer_loc)           # float32 bit patterns


@pytest.mark.parametrize("k,edge,center", [(12, 0.4, 0.4), (16, 0.4, 0.4), (12, 0.25, 0.6), (8, 0.5, 1.0)])
def test_stage_pairs_crp177(hovl_ctx, oracle, crp177, k, edge, center):
    """a3+a4: PairData keys and collision counts (KmerTable.scala:57-149)."""
    bases, offs = crp177
    hs, os_ = _same_settings(oracle, k=k, kmer_edge=edge, kmer_center=center)
    keys, counts, ninst = hovl_ctx.stage_pairs(hs, bases, offs)
    r = oracle.run(os_, bases, offs, threads=4, keep_stages=True)
    assert np.array_equal(keys, r.pair_key)
    assert np.array_equal(counts, r.pair_count)
    assert ninst == r.stats["n_pair_instances"]


def test_stage_pairs_chunked_equals_unchunked(oracle, crp177):
    """PairData built from many small expansion chunks + merge == one chunk."""
    import hovl
    bases, offs = crp177
    hs, os_ = _same_settings(oracle, k=12)
    r = oracle.run(os_, bases, offs, threads=4, keep_stages=True)
    with hovl.Context(0) as ctx:
        ctx.set_option("pair_chunk", 10007)
        keys, counts, ninst = ctx.stage_pairs(hs, bases, offs)
        assert np.array_equal(keys, r.pair_key) and np.array_equal(counts, r.pair_count)
        assert ninst == r.stats["n_pair_instances"]
        res = ctx.run(hs, bases, offs)                 # reference order through the chunked path too
    _assert_records_equal(res.records, r.records)


@pytest.mark.parametrize("k,table_log,cap_hint,mode", [(12, 0, 0, 0), (12, 4, 1, 0), (8, 4, 1, 0), (8, 6, 0, 0),
                                                        (12, 0, 0, 1), (8, 0, 0, 1)])
def test_stage_pairs_gather_tiers_and_sort_mode(oracle, crp177, k, table_log, cap_hint, mode):
    """PairData from the per-read gather (shared-memory tables, overflow to bigger tables and to
    a global table, output-buffer retry) and from the expand+sort path: both equal the oracle."""
    import hovl
    bases, offs = crp177
    hs, os_ = _same_settings(oracle, k=k, kmer_edge=0.5, kmer_center=1.0, order=1, id_mode=1)
    r = oracle.run(os_, bases, offs, threads=4, keep_stages=True)
    with hovl.Context(0) as ctx:
        ctx.set_option("pair_mode", mode)
        ctx.set_option("pair_table_log", table_log)
        if cap_hint:
            ctx.set_option("pair_cap_hint", cap_hint)
        keys, counts, ninst = ctx.stage_pairs(hs, bases, offs)
        assert np.array_equal(keys, r.pair_key) and np.array_equal(counts, r.pair_count)
        assert ninst == r.stats["n_pair_instances"]
        res = ctx.run(hs, bases, offs)
    _assert_records_equal(res.records, r.records)


@pytest.mark.parametrize("window_log2,partition", [(4, 1), (9, 1), (13, 1), (23, 1), (23, 0)])
def test_bucket_of_window_partitions(oracle, window_log2, partition):
    """group stage: bucket_of scattered through window partitions (many windows, ragged last window, one
    window) and straight from the scan give the same PairData and records as the oracle; ragged reads."""
    import hovl
    bases, offs = synth_reads(31, 700, 180, 9000, sub=0.01, ragged=70)
    hs, os_ = _same_settings(oracle, k=12, order=1, id_mode=1, max_ignore=200, max_coll=2**31 - 1)
    r = oracle.run(os_, bases, offs, threads=8, keep_stages=True)
    with hovl.Context(0) as ctx:
        ctx.set_option("group_partition", partition)
        ctx.set_option("group_window_log2", window_log2)
        keys, counts, ninst = ctx.stage_pairs(hs, bases, offs)
        assert np.array_equal(keys, r.pair_key) and np.array_equal(counts, r.pair_count)
        assert ninst == r.stats["n_pair_instances"]
        res = ctx.run(hs, bases, offs)
    _assert_records_equal(res.records, r.records)
    assert len(r.records) > 1000


_SORT_CASES = {}


@pytest.mark.parametrize("tma,ballot,early", [(1, 1, 1), (0, 1, 1), (1, 0, 1), (1, 1, 0), (0, 0, 0)])
@pytest.mark.parametrize("ragged", [0, 40])
def test_radix_sort_variants_give_the_same_table(oracle, tma, ballot, early, ragged):
    """The radix passes with and without the TMA-staged tile, with the lane mask from eight ballots or from match.any,
    with the digit counts published before or after the ranking: same PairData, same records as the oracle.  About 100
    tiles per pass with a partial last one (the look-back runs); equal lengths sort 8-byte tuples, ragged ones 12-byte."""
    import hovl
    if ragged not in _SORT_CASES:
        bases, offs = synth_reads(77, 3000, 150, 60000, sub=0.01, ragged=ragged)
        hs, os_ = _same_settings(oracle, k=12, order=1, id_mode=1, max_ignore=150, max_coll=2**31 - 1)
        _SORT_CASES[ragged] = (bases, offs, hs, oracle.run(os_, bases, offs, threads=os.cpu_count() or 8, keep_stages=True))
    bases, offs, hs, r = _SORT_CASES[ragged]
    with hovl.Context(0) as ctx:
        ctx.set_option("sort_tma", tma)
        ctx.set_option("sort_rank_ballot", ballot)
        ctx.set_option("sort_early_counts", early)
        keys, counts, ninst = ctx.stage_pairs(hs, bases, offs)
        assert np.array_equal(keys, r.pair_key) and np.array_equal(counts, r.pair_count)
        assert ninst == r.stats["n_pair_instances"]
        res = ctx.run(hs, bases, offs)
    _assert_records_equal(res.records, r.records)
    assert len(r.records) > 5000


@pytest.mark.parametrize("compact,L,k", [(1, 100, 12), (0, 100, 12), (1, 333, 16), (0, 333, 16), (1, 12, 12), (1, 40, 20)])
def test_compact_tuples_equal_stored_values(oracle, compact, L, k):
    """Read sets of one common length sort 8-byte (code, read-major index) tuples and recompute (read id, loc)
    in the group stage; the 12-byte (code, read id | loc) form gives the same PairData and records.  L == k is
    the NaN case (loc = 0f/0f: in no section); k = 20 hashes only the first 16 bases."""
    import hovl
    bases, offs = synth_reads(L * 7 + k, 600, L, 40 * L, sub=0.01)
    hs, os_ = _same_settings(oracle, k=k, order=1, id_mode=1, max_ignore=L, max_coll=2**31 - 1, min_coll=3)
    r = oracle.run(os_, bases, offs, threads=8, keep_stages=True)
    with hovl.Context(0) as ctx:
        ctx.set_option("compact_tuples", compact)
        keys, counts, ninst = ctx.stage_pairs(hs, bases, offs)
        assert np.array_equal(keys, r.pair_key) and np.array_equal(counts, r.pair_count)
        assert ninst == r.stats["n_pair_instances"]
        res = ctx.run(hs, bases, offs)
    _assert_records_equal(res.records, r.records)
    assert L <= k or len(r.pair_key) > 100


@pytest.mark.parametrize("k", [12, 16])
def test_stage_align_all_candidate_pairs(hovl_ctx, oracle, crp177, k):
    """a7-a9 on every candidate pair of crp177 (count filter off, duds included)."""
    bases, offs = crp177
    hs, os_ = _same_settings(oracle, k=k, min_coll=1, max_coll=2**31 - 1, filter_valid=0)
    r = oracle.run(os_, bases, offs, threads=4, keep_stages=True)
    ref, _ = oracle.align_pairs(os_, bases, offs, r.disp_a, r.disp_b, threads=4)
    got = hovl_ctx.stage_align(hs, bases, offs, r.disp_a, r.disp_b)
    for f in FIELDS:
        if f != "count":
            assert np.array_equal(got[f], ref[f]), f


@pytest.mark.parametrize("L,mi,indel,ragged", [(150, 0.98, 0.0, 0), (400, 0.98, 0.004, 30), (1000, 0.98, 0.004, 30),
                                               (1000, 0.90, 0.01, 200), (2500, 0.97, 0.002, 0), (1200, 0.80, 0.02, 100),
                                               (6000, 0.98, 0.001, 1000)])
def test_stage_align_synthetic_classes(hovl_ctx, oracle, L, mi, indel, ragged):
    """Every (G,D) band class of the DP kernel, ragged read lengths (mixed classes in one call)."""
    n = 40 if L <= 2500 else 16
    bases, offs = synth_reads(L + 7, n, L, 3 * L, sub=0.01, indel=indel, ragged=ragged)
    hs, os_ = _same_settings(oracle, k=12, filter_valid=0, min_identity=mi, max_ignore=100000)
    pa = np.repeat(np.arange(1, n + 1), n).astype(np.uint32)
    pb = np.tile(np.arange(1, n + 1), n).astype(np.uint32)
    ref, _ = oracle.align_pairs(os_, bases, offs, pa, pb, threads=8)
    got = hovl_ctx.stage_align(hs, bases, offs, pa, pb)
    for f in FIELDS:
        if f != "count":
            assert np.array_equal(got[f], ref[f]), f
    assert (ref["flags"] & 2).sum() > 0


def test_stage_align_k5_shape_10kb(hovl_ctx, oracle):
    """K5's read shape: 10 kb reads, 1 % substitutions, k = 16 -> band width 200 (src/BioLibs.scala:619-620), the
    wide-band warp-per-pair kernel with 7 diagonals per lane (dp_wb.cu).  Also with a few indels and ragged lengths
    (widths 181..201: two classes in one call), and the lane-group kernel it replaces (option dp_wide = 0)."""
    import hovl
    for indel, ragged in ((0.0, 0), (0.001, 1000)):
        n = 14
        bases, offs = synth_reads(5, n, 10000, 16000, sub=0.01, indel=indel, ragged=ragged)
        hs, os_ = _same_settings(oracle, k=16, filter_valid=0, max_ignore=10000)
        pa = np.repeat(np.arange(1, n + 1), n).astype(np.uint32)
        pb = np.tile(np.arange(1, n + 1), n).astype(np.uint32)
        ref, _ = oracle.align_pairs(os_, bases, offs, pa, pb, threads=8)
        got = hovl_ctx.stage_align(hs, bases, offs, pa, pb)
        for f in FIELDS:
            if f != "count":
                assert np.array_equal(got[f], ref[f]), f
        assert (ref["flags"] & 2).sum() > 10
        with hovl.Context(0) as ctx:
            ctx.set_option("dp_wide", 0)
            old = ctx.stage_align(hs, bases, offs, pa, pb)
        for f in FIELDS:
            assert np.array_equal(got[f], old[f]), f


@pytest.mark.parametrize("go,ge,cm", [(-1, -1, (2, -1)), (0, 0, (1, -1)), (-200, -20, None), (-100, -10, (50000, -60000))])
def test_stage_align_wide_band_ties_and_big_costs(hovl_ctx, oracle, go, ge, cm):
    """Wide bands with tie-rich scoring (priority bits, row-major arg-max keys) and with costs too large for the
    25-bit scores of the direction-code kernel (wb::fits fails -> the lane-group kernel takes the pairs)."""
    n = 18
    bases, offs = synth_reads(7, n, 700, 1500, sub=0.05, indel=0.03, ragged=150)
    hs, os_ = _same_settings(oracle, k=12, filter_valid=0, min_identity=0.9, max_ignore=100000)     # widths 56..71
    for st in (hs, os_):
        st.gap_open, st.gap_extend = go, ge
        if cm:
            for i in range(4):
                for j in range(4):
                    st.cost[i][j] = cm[0] if i == j else cm[1]
    pa = np.repeat(np.arange(1, n + 1), n).astype(np.uint32)
    pb = np.tile(np.arange(1, n + 1), n).astype(np.uint32)
    ref, _ = oracle.align_pairs(os_, bases, offs, pa, pb, threads=8)
    got = hovl_ctx.stage_align(hs, bases, offs, pa, pb)
    for f in FIELDS:
        if f != "count":
            assert np.array_equal(got[f], ref[f]), f


def test_stage_align_costs_too_large_for_packed_cells(hovl_ctx, oracle):
    """Narrow bands with costs whose (score << 2) + priority does not fit 32 bits: the reference accepts any Int costs
    (src/BioLibs.scala:119-161), the lane-group kernel with plain 32-bit cells takes those pairs."""
    n = 30
    bases, offs = synth_reads(23, n, 400, 1000, sub=0.01, indel=0.004, ragged=30)
    hs, os_ = _same_settings(oracle, k=12, filter_valid=0, max_ignore=100000)
    for st in (hs, os_):
        st.gap_open, st.gap_extend = -4000000, -500000
        for i in range(4):
            for j in range(4):
                st.cost[i][j] = 3000000 if i == j else -2000000
    pa = np.repeat(np.arange(1, n + 1), n).astype(np.uint32)
    pb = np.tile(np.arange(1, n + 1), n).astype(np.uint32)
    ref, _ = oracle.align_pairs(os_, bases, offs, pa, pb, threads=8)
    got = hovl_ctx.stage_align(hs, bases, offs, pa, pb)
    for f in FIELDS:
        if f != "count":
            assert np.array_equal(got[f], ref[f]), f
    assert (ref["flags"] & 2).sum() > 0


@pytest.mark.parametrize("L,mi,indel,ragged,cm,go,ge,pack", [
    (1000, 0.98, 0.004, 30, (5, -4), -8, -2, 1), (700, 0.97, 0.01, 300, (1, -1), -2, -1, 1),
    (1590, 0.98, 0.004, 400, (2, -3), -5, -2, 1), (130, 0.95, 0.02, 100, (50, -60), -100, -10, 1),
    (1000, 0.98, 0.004, 30, None, -200, -20, 1), (1000, 0.98, 0.004, 30, None, -200, -20, 0),
    (1000, 0.90, 0.01, 200, (1, -1), -1, -1, 1), (800, 0.98, 0.004, 300, (200, -100), -300, -30, 1)])
def test_stage_align_two_pairs_per_thread(oracle, L, mi, indel, ragged, cm, go, ge, pack):
    """16-bit half-word cells (csrc/dp_tp2.cuh): phase 1 alone (HOXD costs), both phases (small
    costs), neither (costs too large / option off) all give the oracle's records."""
    import hovl
    n = 41
    bases, offs = synth_reads(L + 11, n, L, 3 * L, sub=0.01, indel=indel, ragged=ragged)
    hs, os_ = _same_settings(oracle, k=12, filter_valid=0, min_identity=mi, max_ignore=100000)
    for st in (hs, os_):
        st.gap_open, st.gap_extend = go, ge
        if cm:
            for i in range(4):
                for j in range(4):
                    st.cost[i][j] = cm[0] if i == j else cm[1]
    pa = np.repeat(np.arange(1, n + 1), n).astype(np.uint32)
    pb = np.tile(np.arange(1, n + 1), n).astype(np.uint32)
    ref, _ = oracle.align_pairs(os_, bases, offs, pa, pb, threads=8)
    with hovl.Context(0) as ctx:
        ctx.set_option("dp_pack16", pack)
        got = ctx.stage_align(hs, bases, offs, pa, pb)
    for f in FIELDS:
        if f != "count":
            assert np.array_equal(got[f], ref[f]), f
    assert (ref["flags"] & 2).sum() > 0


# ------------------------------------------------------------------ whole path, reference order
@pytest.mark.parametrize("tag,kw", [("crp177_oracle_k12", {}), ("crp177_oracle_k16", {"k": 16})])
def test_run_crp177_byte_identical_ovl(hovl_ctx, oracle, crp177, tag, kw):
    """K2 (crp177, k=16, HOXD1 == built-in matrix) and the default k=12: identical records in the
    reference's (Trove) order, identical .ovl bytes — against the committed vectors and the live oracle."""
    from hovl import io as hio
    bases, offs = crp177
    if "k" in kw:
        kw = dict(kw, cost=hio.read_hoxd(os.path.join(GOLDEN, "HOXD1.txt")))
    hs, os_ = _same_settings(oracle, **kw)
    res = hovl_ctx.run(hs, bases, offs)
    ref = oracle.run(os_, bases, offs, threads=4)
    _assert_records_equal(res.records, ref.records)
    g = np.load(os.path.join(GOLDEN, tag + ".npz"))
    _assert_records_equal(res.records, g["records"].view(oracle.RECORD_DTYPE).reshape(-1))
    assert res.ovl() == open(os.path.join(GOLDEN, tag + ".ovl"), "rb").read()
    for key in ("n_kmers", "n_buckets", "max_bucket", "n_pair_instances", "n_distinct_pairs", "n_dispatched",
                "n_leads", "n_duds", "n_align_valid", "n_ovl_valid", "dp_cells", "n_records"):
        assert res.stats[key] == ref.stats[key], key
    assert res.stats["gpu_launches"] > 10


@pytest.mark.parametrize("n,kind", [(0, 0), (1, 0), (11, 0), (12, 0), (13, 1), (24, 2), (1000, 0), (50000, 1), (300000, 2), (3000000, 0)])
def test_trove_order_on_the_device_equals_the_host_replay(hovl_ctx, n, kind):
    """csrc/trove.cu (all keys claim their probe-path slot at once, rehashes as further rounds) against csrc/order.cpp
    (one insert at a time): random ints, dense small ints, and pair keys (fst << 16) ^ snd (src/KmerTable.scala:73)."""
    import hovl
    rng = np.random.default_rng(n + kind)
    if kind == 0:
        keys = np.unique(rng.integers(-2**31, 2**31, int(n * 1.1) + 8, dtype=np.int64).astype(np.int32))
    elif kind == 1:
        keys = np.arange(4 * n + 5, dtype=np.int32)
    else:
        a, b = rng.integers(1, 30000, 2 * n + 8), rng.integers(1, 30000, 2 * n + 8)
        keys = np.unique(((a << 16) ^ b).astype(np.int64).astype(np.int32))
    keys = keys[rng.permutation(len(keys))][:n]
    assert len(keys) == n
    want = hovl.trove_order(keys)
    got = hovl_ctx.trove_order(keys)
    assert np.array_equal(got, want)


def test_run_crp177_unfiltered_and_canonical(hovl_ctx, oracle, crp177):
    bases, offs = crp177
    hs, os_ = _same_settings(oracle, filter_valid=0)
    _assert_records_equal(hovl_ctx.run(hs, bases, offs).records, oracle.run(os_, bases, offs).records)
    hs, os_ = _same_settings(oracle, order=1, id_mode=1)
    got, ref = hovl_ctx.run(hs, bases, offs).records, oracle.run(os_, bases, offs).records
    _assert_records_equal(got, ref)
    key = got["a"].astype(np.int64) << 32 | got["b"]
    assert (np.diff(key) > 0).all()


def test_run_c_ruddii_full(hovl_ctx, oracle):
    """K1: 32,000 x 100 bp, defaults (k=12): 76.8 M pair instances, 27.4 M distinct pairs, 1.32 M
    dispatched, 379,287 overlaps — identical records in identical order, identical .ovl hash."""
    bases, offs = _c_ruddii_reads()
    hs, _ = _same_settings(oracle)
    res = hovl_ctx.run(hs, bases, offs)
    g = np.load(os.path.join(GOLDEN, "c_ruddii_oracle_k12.npz"))
    ref = g["records"].view(oracle.RECORD_DTYPE).reshape(-1)
    _assert_records_equal(res.records, ref)
    assert hashlib.sha256(res.ovl()).digest() == bytes(g["ovl_sha256"])
    names = oracle.STATS_FIELDS[:14]
    for name, v in zip(names, g["stats"].tolist()):
        assert res.stats[name] == v, name
    keys, counts, _ = hovl_ctx.stage_pairs(hs, bases, offs)
    assert hashlib.sha256(keys.tobytes() + counts.tobytes()).digest() == bytes(g["pair_sha256"])


# ------------------------------------------------------------------ wide ids / synthetic shapes
@pytest.mark.parametrize("n,L,glen,k,kw", [
    (3000, 150, 40000, 12, {}),                                         # K4 shape (short reads, k=12), scaled
    (1500, 1000, 60000, 16, {"max_ignore": 1000, "max_coll": 2**31 - 1}),  # K3 shape (1 kb, 1 % errors), scaled
    (40000, 100, 400000, 12, {}),                                       # ids > 32767: wide keys
])
def test_run_synthetic_canonical_order(hovl_ctx, oracle, n, L, glen, k, kw):
    bases, offs = synth_reads(n + L, n, L, glen, sub=0.01)
    hs, os_ = _same_settings(oracle, k=k, order=1, id_mode=1, **kw)
    res = hovl_ctx.run(hs, bases, offs)
    ref = oracle.run(os_, bases, offs, threads=8)
    _assert_records_equal(res.records, ref.records)
    assert res.ovl() == oracle.format_ovl(ref.records)
    for key in ("n_kmers", "n_buckets", "max_bucket", "n_pair_instances", "n_distinct_pairs", "n_dispatched",
                "n_leads", "n_duds", "n_align_valid", "n_ovl_valid", "dp_cells"):
        assert res.stats[key] == ref.stats[key], key
    assert res.stats["n_ovl_valid"] > 0


def test_run_k5_shape_2000_reads(hovl_ctx, oracle):
    """K5 scaled to 2,000 x 10 kb reads over 2 Mbp (10x), k = 16, wide flags: the whole path with band width 200."""
    n, L = 2000, 10000
    bases, offs = synth_reads(31, n, L, 2_000_000, sub=0.01)
    hs, os_ = _same_settings(oracle, k=16, order=1, id_mode=1, max_ignore=L, max_coll=2**31 - 1)
    res = hovl_ctx.run(hs, bases, offs)
    ref = oracle.run(os_, bases, offs, threads=os.cpu_count() or 8)
    _assert_records_equal(res.records, ref.records)
    assert res.ovl() == oracle.format_ovl(ref.records)
    for key in ("n_kmers", "n_buckets", "n_pair_instances", "n_distinct_pairs", "n_dispatched", "n_duds",
                "n_align_valid", "n_ovl_valid", "dp_cells"):
        assert res.stats[key] == ref.stats[key], key
    assert res.stats["n_ovl_valid"] > 5000


def test_run_device_resident_inputs(hovl_ctx, oracle, crp177):
    import torch
    bases, offs = crp177
    hs, os_ = _same_settings(oracle)
    db = torch.from_numpy(bases).cuda()
    do = torch.from_numpy(offs.view(np.int64)).cuda()
    torch.cuda.synchronize()
    res = hovl_ctx.run_device(hs, db.data_ptr(), do.data_ptr(), len(offs) - 1, len(bases))
    _assert_records_equal(res.records, oracle.run(os_, bases, offs).records)


# ------------------------------------------------------------------ edge cases
def test_edge_cases(hovl_ctx, oracle):
    import hovl
    hs, os_ = _same_settings(oracle)
    # empty input, one read, reads shorter than k, a read of exactly k bases (loc = NaN)
    for reads in ([], [b"ACGTACGTACGTACGTACGT"], [b"ACG", b"", b"ACGTACGTACGT", b"ACGTACGTACGTA"]):
        offs = np.zeros(len(reads) + 1, np.uint64)
        offs[1:] = np.cumsum([len(r) for r in reads])
        bases = np.frombuffer(b"".join(reads), np.uint8)
        res = hovl_ctx.run(hs, bases, offs)
        ref = oracle.run(os_, bases, offs)
        _assert_records_equal(res.records, ref.records)
        assert res.stats["n_kmers"] == ref.stats["n_kmers"]
        if len(reads):
            c, r, l = hovl_ctx.stage_table(hs, bases, offs)
            rr = oracle.run(os_, bases, offs, keep_stages=True)
            assert np.array_equal(c, rr.kmer_code) and np.array_equal(r, rr.kmer_rid) and np.array_equal(l, rr.kmer_loc)
    # > 32767 reads need the wide key
    offs = (np.arange(32769 + 1) * 12).astype(np.uint64)
    bases = np.frombuffer(b"ACGTACGTACGT" * 32769, np.uint8)
    with pytest.raises(hovl.HovlError) as e:
        hovl_ctx.run(hs, bases, offs)
    assert e.value.code == hovl.E_UNSUPPORTED
    with pytest.raises(hovl.HovlError):
        hovl_ctx.run(hovl.default_settings(gap_open=0, k=0), bases[:12], offs[:2])


def test_bad_base_and_short_read_errors(hovl_ctx, oracle, crp177):
    """Non-ACGT bytes hash as 'A' in the k-mer code (ObjectStore.scala:60-63) but cannot be aligned
    (MatchError in costFunc); a B shorter than the band cannot be aligned either."""
    import hovl
    bases, offs = crp177
    b2 = bases.copy()
    b2[int(offs[10]) + 50] = ord("N")
    hs, os_ = _same_settings(oracle, filter_valid=0)
    res = hovl_ctx.run(hs, b2, offs, strict=False)
    ref = oracle.run(os_, b2, offs)
    assert res.rc == ref.rc == hovl.E_BAD_BASE
    _assert_records_equal(res.records, ref.records)
    assert (res.records["flags"] & hovl.F_BAD_BASE).sum() > 0
    with pytest.raises(hovl.HovlError):
        hovl_ctx.run(hs, b2, offs)
    # short B: identity 0.5 -> width 51 for A of 100 bases, B of 40 bases
    bs, os2 = synth_reads(3, 30, 100, 300, sub=0.0)
    keep = np.ones(len(bs), bool)
    keep[int(os2[3]) + 40:int(os2[4])] = False
    lens = np.diff(os2.astype(np.int64)); lens[3] = 40
    bs = bs[keep]; os3 = np.zeros(31, np.uint64); os3[1:] = np.cumsum(lens)
    hs, os_ = _same_settings(oracle, filter_valid=0, min_identity=0.5, min_coll=1)
    res = hovl_ctx.run(hs, bs, os3, strict=False)
    ref = oracle.run(os_, bs, os3)
    assert res.rc == ref.rc
    _assert_records_equal(res.records, ref.records)

"""Pins the CPU oracle (oracle/hovl_oracle.c) against what the reference ships for this path
(SURVEY.md §8c): the AMOS golden overlap sets as superset pins, the reference's own
"no pair dispatched twice" invariant, the committed oracle vectors, and the stage counts."""
import hashlib
import os

import numpy as np
import pytest

from conftest import GOLDEN


def _ovl_set(arr):
    """AMOS lists each overlapping pair once, lower read id first; an overlap (a,b,ahg,bhg) with
    a > b is the same overlap as (b,a,-ahg,-bhg)."""
    arr = np.asarray(arr, np.int64).reshape(-1, 4).copy()
    sw = arr[:, 0] > arr[:, 1]
    arr[sw] = np.stack([arr[sw, 1], arr[sw, 0], -arr[sw, 2], -arr[sw, 3]], 1)
    return set(map(tuple, arr.tolist()))


def _parse_ovl(path):
    import re
    txt = open(path).read()
    recs = re.findall(r"rds:(\d+),(\d+)\s+scr:(-?\d+)\s+ahg:(-?\d+)\s+bhg:(-?\d+)", txt)
    return np.array([[int(a), int(b), int(c), int(d)] for a, b, _, c, d in recs], np.int64)


def test_crp177_defaults_subset_of_amos_golden(oracle, crp177):
    bases, offs = crp177
    r = oracle.run(oracle.default_settings(), bases, offs, threads=4, keep_stages=True)
    assert r.rc == 0
    st = r.stats
    # work sizes found when the reference's semantics were surveyed
    assert (st["n_kmers"], st["n_buckets"], st["max_bucket"]) == (19847, 1164, 57)
    assert (st["n_distinct_pairs"], st["n_dispatched"], st["n_leads"]) == (5412, 2452, 222)
    assert (st["n_duds"], st["n_align_valid"], st["n_ovl_valid"]) == (37, 2388, 2388)
    gold = _ovl_set(_parse_ovl(os.path.join(GOLDEN, "crp177.ovl")))
    assert len(gold) == 2598
    mine = _ovl_set(np.stack([r.records[f] for f in ("a", "b", "ahg", "bhg")], 1))
    assert len(mine) == 2388 and mine <= gold          # same pairs, same hangs
    # no pair dispatched twice (src/Project4.scala:381-384, :399-402)
    pairs = set(zip(r.disp_a.tolist(), r.disp_b.tolist()))
    assert len(pairs) == len(r.disp_a)


def test_crp177_k16_hoxd1(oracle, crp177):
    from hovl import io as hio
    bases, offs = crp177
    hoxd = hio.read_hoxd(os.path.join(GOLDEN, "HOXD1.txt"))
    s = oracle.default_settings(k=16, cost=hoxd)
    d = oracle.default_settings(k=16)
    assert [list(r) for r in s.cost] == [list(r) for r in d.cost]      # HOXD1.txt == defaultHOXD
    r = oracle.run(s, bases, offs, threads=4)
    assert (r.stats["n_kmers"], r.stats["n_buckets"], r.stats["max_bucket"]) == (18955, 1194, 18)
    assert r.stats["n_distinct_pairs"] == 2387 and r.stats["n_dispatched"] == 2387
    gold = _ovl_set(_parse_ovl(os.path.join(GOLDEN, "crp177.ovl")))
    mine = _ovl_set(np.stack([r.records[f] for f in ("a", "b", "ahg", "bhg")], 1))
    assert mine <= gold


def test_hoxd2_pair_list_form_is_rejected_like_the_reference():
    from hovl import io as hio
    with pytest.raises(ValueError):
        hio.read_hoxd(os.path.join(GOLDEN, "HOXD2.txt"))


@pytest.mark.parametrize("tag,kw", [("crp177_oracle_k12", {}), ("crp177_oracle_k16", {"k": 16})])
def test_committed_vectors_match_live_oracle(oracle, crp177, tag, kw):
    bases, offs = crp177
    g = np.load(os.path.join(GOLDEN, tag + ".npz"))
    r = oracle.run(oracle.default_settings(**kw), bases, offs, threads=2, keep_stages=True)
    assert np.array_equal(r.records, g["records"].view(oracle.RECORD_DTYPE).reshape(-1))
    assert np.array_equal(r.pair_key, g["pair_key"]) and np.array_equal(r.pair_count, g["pair_count"])
    ovl = oracle.format_ovl(r.records)
    assert ovl == open(os.path.join(GOLDEN, tag + ".ovl"), "rb").read()
    assert hashlib.sha256(ovl).digest() == bytes(g["ovl_sha256"])


def test_c_ruddii_vectors_subset_of_amos_golden(oracle):
    """32,000 x 100 bp reads rebuilt from the bank layout; the oracle's 379,287 overlaps (committed
    vector) must all be among AMOS's 388,898 with the same hangs."""
    g = np.load(os.path.join(GOLDEN, "c_ruddii_oracle_k12.npz"))
    recs = g["records"].view(oracle.RECORD_DTYPE).reshape(-1)
    amos = np.load(os.path.join(GOLDEN, "c_ruddii_amos_ovl.npz"))["ovl"]
    assert len(amos) == 388898
    gold = _ovl_set(amos)
    mine = _ovl_set(np.stack([recs[f] for f in ("a", "b", "ahg", "bhg")], 1))
    assert len(recs) == 379287 and len(mine) > 180000     # (a,b) and (b,a) may both be emitted
    assert mine <= gold
    assert len(set(zip(recs["a"].tolist(), recs["b"].tolist()))) == len(recs)
    # the reads were cut from the contig at known offsets: every emitted ahg is the true offset
    lay = np.load(os.path.join(GOLDEN, "c_ruddii_layout.npy")).astype(np.int64)
    assert np.array_equal(lay[recs["b"] - 1] - lay[recs["a"] - 1], recs["ahg"])
    assert np.array_equal(recs["ahg"], recs["bhg"])


def test_trove_model_basics(oracle):
    """App. B of SURVEY.md: capacity 23 at birth, rehash to nextPrime(2*cap) past load 0.5,
    iteration from the top slot down, identity hash."""
    import ctypes as C
    L = oracle.lib()
    m = L.trove_new()
    assert L.trove_capacity(m) == 23
    isnew = C.c_int()
    for k in range(11):
        L.trove_put_if_absent(m, k, C.byref(isnew))
    assert L.trove_capacity(m) == 23 and L.trove_size(m) == 11
    L.trove_put_if_absent(m, 11, C.byref(isnew))           # 12 > maxSize 11 -> 47
    assert L.trove_capacity(m) == 47
    keys = (C.c_int32 * 12)()
    L.trove_iterate(m, keys, None)
    assert list(keys) == list(range(11, -1, -1))            # slot = key % 47, walked downwards
    L.trove_free(m)
    # collisions: 1 and 24 share slot 1 in a 23-slot table; probe = 1 + h % 21
    m = L.trove_new()
    for k in (1, 24, -5):
        L.trove_put_if_absent(m, k, C.byref(isnew))
    keys = (C.c_int32 * 3)()
    L.trove_iterate(m, keys, None)
    # 24 -> slot 1 taken -> probe 1+24%21=4 -> slot 1-4+23=20 ; -5 & 0x7fffffff = 2147483643 -> %23
    # h = 2^31-5 -> h % 23 = 1 (taken) -> probe 1 + h % 21 = 19 -> slot 1-19+23 = 5
    h = (-5) & 0x7fffffff
    assert h % 23 == 1 and 1 + h % 21 == 19
    slots = {1: 1, 24: 20, -5: 5}
    assert list(keys) == [k for k, _ in sorted(slots.items(), key=lambda kv: -kv[1])]
    L.trove_free(m)


def test_readers(oracle, tmp_path):
    from hovl import io as hio
    p = tmp_path / "x.fa"
    p.write_bytes(b">r1 something\nacgt\nNNAC\n>r2\n\n>r3\nTT")
    b, o = hio.read_seq(str(p))
    b2, o2 = oracle.read_fasta(str(p))
    assert bytes(b) == b"ACGTNNACTT" == bytes(b2)
    assert o.tolist() == [0, 8, 8, 10] == o2.tolist()
    bad = tmp_path / "bad.fa"
    bad.write_bytes(b"ACGT\n")
    with pytest.raises(ValueError):
        hio.read_seq(str(bad))
    with pytest.raises(ValueError):
        oracle.read_fasta(str(bad))
    assert hio.read_hoxd(os.path.join(GOLDEN, "HOXD1.txt")) == oracle.read_hoxd(os.path.join(GOLDEN, "HOXD1.txt"))

"""World-size-N gloo worker for tests/test_dist_cpu.py: drives hovl.dist.ShardedOverlapper (the
same exchange code the GPU path uses) with the CPU oracle standing in for the three device
stages, and checks the gathered records against a single-process oracle run.  Launched with
torch.distributed.run; test infrastructure, not product code."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "sequence-aligner_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import hovl                                   # noqa: E402  (ctypes mirror; loads without a GPU)
from hovl import dist as hdist                # noqa: E402
from oracle import oracle as O                # noqa: E402
from conftest import synth_reads              # noqa: E402


def mix32(x):
    x = x.astype(np.uint64)
    x ^= x >> 16; x = (x * 0x85ebca6b) & 0xffffffff
    x ^= x >> 13; x = (x * 0xc2b2ae35) & 0xffffffff
    x ^= x >> 16
    return x


def owner_of_code(code, parts):
    return ((mix32(code) * parts) >> 32).astype(np.int64)


def owner_of_code_range(code, k, parts):
    return ((code.astype(np.uint64) * parts) >> np.uint64(2 * min(k, 16))).astype(np.int64)


def owner_of_read(rid, n_total, parts):
    return ((rid.astype(np.uint64) - 1) * parts // max(n_total, 1)).astype(np.int64)


def f32(bits):
    return np.asarray(bits, np.uint32).view(np.float32)


class OracleStages:
    """Same three stages as hovl.dist.GpuStages, computed on the CPU from the oracle."""

    def __init__(self, settings):
        self.s = settings
        self.zero = {k: 0 for k in hdist.STAT_SUM + hdist.STAT_MAX}
        self.by_range = False            # set by ShardedOverlapper(mode="replicate"): ranks own code ranges

    def encode(self, d_bases, d_offsets, lo, n_total, parts):
        bases = d_bases.numpy()
        offs = d_offsets.numpy().astype(np.uint64)
        r = O.run(self.s, bases, offs, threads=1, keep_stages=True)
        code, rid, loc = r.kmer_code, r.kmer_rid.astype(np.uint64) + lo, r.kmer_loc
        dest = owner_of_code_range(code, self.s.k, parts) if self.by_range else owner_of_code(code, parts)
        order = np.argsort(dest, kind="stable")
        counts = np.bincount(dest, minlength=parts).tolist()
        vals = ((rid << np.uint64(32)) | loc.astype(np.uint64))[order].view(np.int64)
        lens = (offs[1:] - offs[:-1]).astype(np.int32)
        self.zero["n_kmers"] = len(code)
        return (torch.from_numpy(code[order].view(np.int32).copy()), torch.from_numpy(vals.copy()), counts,
                torch.from_numpy(bases.astype(np.int32)), torch.from_numpy(lens), torch.zeros(len(lens), dtype=torch.int32))

    def _count(self, code, vals, fst_lo=None, fst_hi=None):
        """PairData of the given tuples: {(fst, snd): count} and the number of instances (optionally only fst in a range)."""
        s = self.s
        rid, loc = (vals >> np.uint64(32)).astype(np.int64), f32((vals & np.uint64(0xffffffff)).astype(np.uint32))
        head, tail, ml, mt = (f32([b])[0] for b in (s.head_f32, s.tail_f32, s.mid_lead_f32, s.mid_tail_f32))
        acc = {}
        order = np.argsort(code, kind="stable")
        cs = code[order]
        cuts = np.flatnonzero(np.r_[True, cs[1:] != cs[:-1], True])
        ninst = 0
        for lo_, hi_ in zip(cuts[:-1], cuts[1:]):
            idx = order[lo_:hi_]
            rr, ll = rid[idx], loc[idx]
            md = [(int(a), float(b)) for a, b in zip(rr, ll) if ml <= b <= mt]
            for sel in (ll <= head, tail <= ll):                     # KmerTable.scala:118-128
                for er, el in zip(rr[sel], ll[sel]):
                    for mr, mloc in md:
                        if er == mr:
                            continue
                        fst, snd = (int(er), mr) if el > mloc else (mr, int(er))   # :63-71
                        if fst_lo is not None and not (fst_lo <= fst <= fst_hi):
                            continue
                        acc[(fst, snd)] = acc.get((fst, snd), 0) + 1
                        ninst += 1
        return acc, ninst

    def sort(self, r_codes, r_vals):
        """Replicated-table variant: stable sort of the received code range, in place."""
        code = r_codes.numpy().view(np.uint32)
        order = np.argsort(code, kind="stable")
        r_codes.copy_(torch.from_numpy(code[order].view(np.int32).copy()))
        r_vals.copy_(torch.from_numpy(r_vals.numpy()[order].copy()))

    def pairs_local(self, all_codes, all_vals, all_packed, all_len, all_bad, n_total, lo, n_local):
        code = all_codes.numpy().view(np.uint32)
        assert (np.diff(code.astype(np.int64)) >= 0).all(), "the gathered slices must concatenate into the sorted table"
        acc, ninst = self._count(code, all_vals.numpy().view(np.uint64), lo + 1, lo + n_local)
        self.zero["n_pair_instances"] = ninst
        id_bits = max(1, int(n_total).bit_length())
        items = sorted(acc.items())
        keys = torch.from_numpy(np.array([(f << id_bits) | b for (f, b), _ in items], np.uint64).view(np.int64).copy())
        cnts = torch.from_numpy(np.array([c for _, c in items], np.int32))
        recs, st = self.align(keys, cnts, all_packed, all_len, all_bad, n_total)
        return recs, st, None

    def pairs(self, r_codes, r_vals, n_total, parts):
        acc, ninst = self._count(r_codes.numpy().view(np.uint32), r_vals.numpy().view(np.uint64))
        id_bits = max(1, int(n_total).bit_length())
        keys = np.array([(f << id_bits) | b for f, b in acc], np.uint64)
        cnts = np.array(list(acc.values()), np.int32)
        dest = owner_of_read(keys >> np.uint64(id_bits), n_total, parts) if len(keys) else np.zeros(0, np.int64)
        o = np.argsort(dest, kind="stable")
        self.zero["n_pair_instances"] = ninst
        return (torch.from_numpy(keys[o].view(np.int64).copy()), torch.from_numpy(cnts[o].copy()),
                np.bincount(dest, minlength=parts).tolist())

    def align(self, r_keys, r_cnts, all_packed, all_len, all_bad, n_total):
        s = self.s
        keys = r_keys.numpy().view(np.uint64)
        uk, inv = np.unique(keys, return_inverse=True)
        cnt = np.bincount(inv, weights=r_cnts.numpy(), minlength=len(uk)).astype(np.int64)
        keep = (cnt >= s.min_coll) & (cnt <= s.max_coll)
        id_bits = max(1, int(n_total).bit_length())
        a = (uk[keep] >> np.uint64(id_bits)).astype(np.uint32)
        b = (uk[keep] & np.uint64((1 << id_bits) - 1)).astype(np.uint32)
        bases = all_packed.numpy().astype(np.uint8)
        offs = np.r_[0, np.cumsum(all_len.numpy().astype(np.int64))].astype(np.uint64)
        recs, cells = O.align_pairs(s, bases, offs, a, b, threads=2)
        recs["count"] = cnt[keep]
        st = dict(self.zero)
        st.update(n_distinct_pairs=len(uk), n_dispatched=len(a), n_duds=int((recs["flags"] & 1).sum()),
                  n_align_valid=int((recs["flags"] & 2).sum() // 2), n_ovl_valid=int((recs["flags"] & 4).sum() // 4),
                  dp_cells=cells)
        if s.filter_valid:
            recs = recs[(recs["flags"] & 2) != 0]
        st["n_records"] = len(recs)
        return recs, st

    # ---- alignment re-balance by DP cells: the CPU mirror of hovl_shard_select / _split / _fetch_selected / _align_pairs
    def select(self, r_keys, r_cnts, all_len, n_total):
        """Sum equal keys, count filter; keeps (a, b, count) sorted by (a, b) and the exclusive prefix of their DP cost
        lenA * width + (lenA + 1) * (width + 1)  (csrc/rebalance.cu CostIn, src/BioLibs.scala:619-624)."""
        s = self.s
        keys = r_keys.numpy().view(np.uint64)
        uk, inv = np.unique(keys, return_inverse=True)
        cnt = np.bincount(inv, weights=r_cnts.numpy(), minlength=len(uk)).astype(np.int64)
        keep = (cnt >= s.min_coll) & (cnt <= s.max_coll)
        id_bits = max(1, int(n_total).bit_length())
        a = (uk[keep] >> np.uint64(id_bits)).astype(np.int64)
        b = (uk[keep] & np.uint64((1 << id_bits) - 1)).astype(np.int64)
        n = all_len.numpy().astype(np.int64)[a - 1]
        omi = np.float32(1.0) - f32([s.min_identity_f32])[0]
        w = np.maximum(s.k, (n.astype(np.float32) * omi).astype(np.int64) + 1)
        cost = n * w + (n + 1) * (w + 1)
        self._sel = (a.astype(np.int32), b.astype(np.int32), cnt[keep].astype(np.int32))
        self._excl = np.r_[0, np.cumsum(cost)].astype(np.int64)
        self._ndistinct = len(uk)
        return len(a), int(self._excl[-1])

    def split(self, cost_before, cost_total, parts):
        """Pair i sits at cost_before + excl[i] of the job's cost line and goes to rank floor(position * parts / total)."""
        n = len(self._sel[0])
        cut = [0] * (parts + 1)
        cut[parts] = n
        for p in range(1, parts):
            thr = -(-p * cost_total // parts)                     # ceil(p * total / parts)
            cut[p] = int(np.searchsorted(cost_before + self._excl[:n], thr, side="left"))
        return [cut[p + 1] - cut[p] for p in range(parts)]

    def fetch_selected(self, n_pairs, dev):
        a, b, c = self._sel
        assert n_pairs == len(a)
        return torch.from_numpy(np.stack([a, b, c], axis=1).reshape(-1).copy())

    def align_pairs(self, abc, all_packed, all_len, all_bad, n_total):
        s = self.s
        t = abc.numpy().reshape(-1, 3)
        a, b, cnt = t[:, 0].astype(np.uint32), t[:, 1].astype(np.uint32), t[:, 2].astype(np.int64)
        if len(a) > 1:                                                # what a rank receives is sorted by (a, b)
            key = a.astype(np.int64) << 32 | b.astype(np.int64)
            assert (np.diff(key) > 0).all(), "re-balanced pairs must arrive in canonical order"
        bases = all_packed.numpy().astype(np.uint8)
        offs = np.r_[0, np.cumsum(all_len.numpy().astype(np.int64))].astype(np.uint64)
        recs, cells = O.align_pairs(s, bases, offs, a, b, threads=2)
        recs["count"] = cnt
        st = dict(self.zero)
        st.update(n_distinct_pairs=getattr(self, "_ndistinct", 0), n_dispatched=len(a), n_duds=int((recs["flags"] & 1).sum()),
                  n_align_valid=int((recs["flags"] & 2).sum() // 2), n_ovl_valid=int((recs["flags"] & 4).sum() // 4),
                  dp_cells=cells)
        if s.filter_valid:
            recs = recs[(recs["flags"] & 2) != 0]
        st["n_records"] = len(recs)
        return recs, st

    def kernel_times(self):
        return {}


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    # owner functions of the test == the library's (host-only exports of libhovl.so)
    L = hovl.lib()
    for c in (0, 1, 12345, 0xffffffff, 0x9e3779b9):
        assert int(owner_of_code(np.array([c], np.uint32), world)[0]) == L.hovl_shard_of_code(c, world)
        for k in (12, 16, 20):
            cc = c & ((1 << (2 * min(k, 16))) - 1)
            assert int(owner_of_code_range(np.array([cc], np.uint32), k, world)[0]) == L.hovl_shard_of_code_range(cc, k, world)
    for case, (n, rl, glen, k, ragged) in enumerate([(120, 100, 900, 12, 0), (90, 140, 700, 16, 30), (96, 0, 900, 12, 0)]):
        skewed = rl == 0
        if skewed:
            # the second half of the reads is three times as long: the ranks that own them carry most of the DP cells, so
            # the re-balance by cost has to move pairs (and the records must not change)
            b1, o1 = synth_reads(300, n // 2, 100, glen, sub=0.01)
            b2, o2 = synth_reads(301, n - n // 2, 300, glen, sub=0.01)
            bases = np.concatenate([b1, b2])
            offs = np.concatenate([o1, o2[1:] + o1[-1]]).astype(np.uint64)
        else:
            bases, offs = synth_reads(100 + case, n, rl, glen, sub=0.01, ragged=ragged)
        for rid in (1, n // 2, n):
            assert int(owner_of_read(np.array([rid]), n, world)[0]) == L.hovl_shard_of_read(rid, n, world)
        s = O.default_settings(k=k, order=1, id_mode=1, max_ignore=1000, max_coll=2**31 - 1)
        lo, hi = hdist.read_slice(n, rank, world)
        b0, b1 = int(offs[lo]), int(offs[hi])
        local = torch.from_numpy(bases[b0:b1].copy())
        loffs = torch.from_numpy((offs[lo:hi + 1] - offs[lo]).astype(np.int64))
        ref = O.run(s, bases, offs, threads=2)
        for mode, rebalance in (("partials", True), ("partials", False), ("replicate", True)):
            eng = hdist.ShardedOverlapper(None, s, stages=OracleStages(s), mode=mode, rebalance=rebalance)
            res = eng.run_device(local, lo, hi, n, d_offsets=loffs)
            if mode == "partials":
                assert (eng.last_rebalance is not None) == rebalance
                if rebalance and skewed and world > 1:
                    assert eng.last_rebalance["moved"], eng.last_rebalance
                    assert max(eng.last_rebalance["cost"]) * world > 1.08 * sum(eng.last_rebalance["cost"])
            if rank == 0:
                assert len(res.records) == len(ref.records) > 0, (mode, len(res.records), len(ref.records))
                for f in ref.records.dtype.names:
                    assert np.array_equal(res.records[f], ref.records[f]), (mode, f)
                for key in ("n_kmers", "n_pair_instances", "n_dispatched", "n_duds", "n_align_valid",
                            "n_ovl_valid", "dp_cells", "n_records") + (("n_distinct_pairs",) if mode == "partials" else ()):
                    assert res.stats[key] == ref.stats[key], (mode, key, res.stats[key], ref.stats[key])
            else:
                assert res.records is None
            # copy=True detaches the records from the buffer the next run reuses
            kept = eng.run_device(local, lo, hi, n, d_offsets=loffs, copy=True)
            again = eng.run_device(local, lo, hi, n, d_offsets=loffs)
            if rank == 0:
                assert kept.records.base is None or not np.shares_memory(kept.records, again.records)
                assert np.array_equal(kept.records, again.records)
    dist.barrier()
    if rank == 0:
        print("DIST_CPU_OK world=%d" % world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

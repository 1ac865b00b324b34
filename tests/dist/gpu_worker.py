"""One rank of the sharded GPU path (tests/test_gpu_dist.py; launched with torch.distributed.run).
Rank 0 checks the gathered records against the single-GPU hovl_run on the same reads and, for
the small case, against the CPU oracle."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "sequence-aligner_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import hovl                                   # noqa: E402
from hovl import dist as hdist                # noqa: E402
from conftest import synth_reads              # noqa: E402

STATS = ("n_kmers", "n_pair_instances", "n_distinct_pairs", "n_dispatched", "n_leads", "n_duds", "n_align_valid",
         "n_ovl_valid", "dp_cells", "n_records")


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    rank, world = dist.get_rank(), dist.get_world_size()
    ctx = hovl.Context(local)
    cases = [(600, 100, 6000, 12, 0, {}, True, False),
             (3000, 150, 40000, 12, 40, {}, False, False),
             (2500, 1000, 90000, 16, 0, {"max_ignore": 1000, "max_coll": 2**31 - 1}, False, False),
             # reads sorted by length (300 .. 1200 bases): the DP cost of a contiguous slice of reads grows with the rank, so
             # the alignment re-balance has to move pairs (csrc/rebalance.cu)
             (2400, 1200, 90000, 16, 900, {"max_ignore": 1200, "max_coll": 2**31 - 1}, False, True)]
    for n, rl, glen, k, ragged, kw, with_oracle, by_length in cases:
        bases, offs = synth_reads(n + rl, n, rl, glen, sub=0.01, ragged=ragged)
        if by_length:
            lens = np.diff(offs.astype(np.int64))
            order = np.argsort(lens, kind="stable")
            bases = np.concatenate([bases[int(offs[i]):int(offs[i + 1])] for i in order])
            offs = np.concatenate([[0], np.cumsum(lens[order])]).astype(np.uint64)
        s = hovl.default_settings(k=k, order=1, id_mode=1, **kw)
        lo, hi = hdist.read_slice(n, rank, world)
        b0, b1 = int(offs[lo]), int(offs[hi])
        d_local = torch.from_numpy(bases[b0:b1].copy()).to(dev)
        d_offs = torch.from_numpy((offs[lo:hi + 1] - offs[lo]).astype(np.int64)).to(dev)
        ref = None
        if rank == 0:
            with hovl.Context(local) as single:
                ref = single.run(s, bases, offs)
        # "partials": hashed k-mer shards + exchange of partial pair counts; "replicate": code ranges, the sorted table
        # gathered on every rank, pairs counted by the owner of fst (include/hovl.h); twice each: buffers are reused
        for mode in ("partials", "replicate", "replicate", "partials"):
            eng = hdist.ShardedOverlapper(ctx, s, mode=mode)
            res = eng.run_device(d_local, lo, hi, n, d_offsets=d_offs)
            if by_length and mode == "partials" and world > 1:
                assert eng.last_rebalance is not None and eng.last_rebalance["moved"], eng.last_rebalance
                cost = eng.last_rebalance["cost"]
                assert max(cost) * world > 1.3 * sum(cost), cost          # the slices really were unequal
                if rank == 0:
                    print("rebalance: DP cost by rank before the move", cost, flush=True)
            if rank == 0:
                assert len(res.records) == len(ref.records) > 0, (mode, len(res.records), len(ref.records))
                for f in ref.records.dtype.names:
                    assert np.array_equal(res.records[f], ref.records[f]), (mode, f)
                for key in STATS + ("n_buckets",):
                    assert res.stats[key] == ref.stats[key], (mode, key, res.stats[key], ref.stats[key])
                if with_oracle:
                    from oracle import oracle as O
                    os_ = O.default_settings(k=k, order=1, id_mode=1, **kw)
                    oref = O.run(os_, bases, offs, threads=4)
                    for f in oref.records.dtype.names:
                        assert np.array_equal(res.records[f], oref.records[f]), (mode, f)
            del res
            eng.close()
            dist.barrier()
    ctx.close()
    if rank == 0:
        print("DIST_GPU_OK world=%d" % world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Alignment re-balance by DP cells on a skewed read set (csrc/rebalance.cu): the same reads, sorted by length so that a
contiguous slice of read ids per rank means unequal DP work, aligned with and without moving pairs.
  python -m torch.distributed.run --nproc-per-node N tools/rebalance_demo.py > profiles/r02_rebalance_Ngpu.json
Prints one JSON line: per-rank DP milliseconds and device-timed step time for both settings, and the records' sha256."""
import hashlib
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "sequence-aligner_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)
import hovl                                   # noqa: E402
from hovl import dist as hdist                # noqa: E402
from conftest import synth_reads              # noqa: E402


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    rank, world = dist.get_rank(), dist.get_world_size()
    n, rl, glen = 60000, 1500, 2_000_000                 # lengths 300 .. 1500, 27 x coverage
    bases, offs = synth_reads(11, n, rl, glen, sub=0.01, ragged=1200)
    lens = np.diff(offs.astype(np.int64))
    order = np.argsort(lens, kind="stable")
    bases = np.concatenate([bases[int(offs[i]):int(offs[i + 1])] for i in order])
    offs = np.concatenate([[0], np.cumsum(lens[order])]).astype(np.uint64)
    s = hovl.default_settings(k=16, order=1, id_mode=1, max_ignore=rl, max_coll=2**31 - 1)
    lo, hi = hdist.read_slice(n, rank, world)
    d_local = torch.from_numpy(bases[int(offs[lo]):int(offs[hi])].copy()).to(dev)
    d_offs = torch.from_numpy((offs[lo:hi + 1] - offs[lo]).astype(np.int64)).to(dev)
    ctx = hovl.Context(local)
    out = {"workload": "%d reads of 300..1500 bp sorted by length, %d bp genome, k=16, wide flags" % (n, glen), "n_gpus": world}
    for name, reb in (("without", False), ("with", True)):
        eng = hdist.ShardedOverlapper(ctx, s, rebalance=reb)
        for _ in range(3):
            res = eng.run_device(d_local, lo, hi, n, d_offsets=d_offs)
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            res = eng.run_device(d_local, lo, hi, n, d_offsets=d_offs)
        e1.record()
        dist.barrier(); torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / 3], device=dev, dtype=torch.float64)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dp = torch.tensor([res.local_stats["ns_dp"] / 1e6], device=dev, dtype=torch.float64)
        dps = [torch.zeros_like(dp) for _ in range(world)]
        dist.all_gather(dps, dp)
        if rank == 0:
            out[name + "_rebalance"] = {"ms_per_step": float(ms.item()), "dp_ms_by_rank": [round(float(x.item()), 3) for x in dps],
                                       "dp_cost_by_rank_before_move": (eng.last_rebalance or {}).get("cost"),
                                       "moved": (eng.last_rebalance or {}).get("moved"),
                                       "records_sha256": hashlib.sha256(np.ascontiguousarray(res.records).tobytes()).hexdigest(),
                                       "n_records": int(len(res.records))}
        del res
        eng.close()
    if rank == 0:
        print(json.dumps(out))
    ctx.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

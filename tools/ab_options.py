#!/usr/bin/env python
"""A/B timing of libhovl options on one GPU, one process, one synthetic read set.

  python tools/ab_options.py --workload K3 --scale 1 --steps 5 --warmup 3 \
      --set "" --set sort_rank_ballot=0 --set sort_tma=0 --set sort_rank_ballot=0,sort_tma=0

For every --set (a comma-separated list of name=value passed to hovl_set_option; "" = the defaults) it runs hovl_run_device
on the same reads already in HBM, times K steps with CUDA events on the library's stream and prints one JSON line: ms per
step, the per-kernel device times of the last step and the sha256 of the records (which must not depend on the options).
Options set by one --set are put back to the values named in --reset before the next one."""
import argparse
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sequence-aligner_b200"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="K3")
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--shape", default="wide")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--set", action="append", default=[], dest="sets")
    ap.add_argument("--reset", default="sort_rank_ballot=1,sort_tma=1,sort_early_counts=1",
                    help="name=value list applied before every --set")
    args = ap.parse_args()
    import torch
    import hovl
    from hovl import synth
    if not torch.cuda.is_available():
        raise SystemExit("needs a CUDA device: libhovl has no CPU fallback")
    dev = torch.device("cuda", 0)
    w = synth.scaled(args.workload, args.scale)
    n, L = w["reads"], w["length"]
    h_bases = torch.empty(n * L, dtype=torch.uint8, pin_memory=True)
    _, offs = synth.generate(w, out=h_bases.numpy())
    h_offs = torch.from_numpy(np.asarray(offs, dtype=np.uint64).view(np.int64).copy())
    settings = hovl.default_settings(**synth.shape_flags(w, args.shape))
    stream = torch.cuda.Stream(device=dev)
    ctx = hovl.Context(0)
    ctx.set_option("stream", stream.cuda_stream)
    ctx.set_option("kernel_timing", 1)
    d_bases, d_offs = h_bases.to(dev), h_offs.to(dev)
    torch.cuda.synchronize()
    n_bases = int(offs[-1])

    def apply(spec):
        for kv in [x for x in spec.split(",") if x]:
            name, value = kv.split("=", 1)
            ctx.set_option(name, int(value))

    first = None
    for spec in (args.sets or [""]):
        apply(args.reset)
        apply(spec)
        res = None
        for _ in range(args.warmup):
            res = ctx.run_device(settings, d_bases.data_ptr(), d_offs.data_ptr(), n, n_bases)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            e0.record(stream)
            for _ in range(args.steps):
                res = ctx.run_device(settings, d_bases.data_ptr(), d_offs.data_ptr(), n, n_bases)
            e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.steps
        sha = hashlib.sha256(np.ascontiguousarray(res.records).tobytes()).hexdigest()
        first = first or sha
        kt = ctx.kernel_times()
        top = sorted(((k, c, ns / 1e6) for k, (c, ns) in kt.items()), key=lambda t: -t[2])[:14]
        print(json.dumps({"options": spec or "(defaults)", "workload": "%s x %g" % (args.workload, args.scale),
                          "ms_per_step": round(ms, 3), "reads_per_s": round(n / (ms / 1e3)),
                          "n_records": int(len(res.records)), "records_sha256": sha[:12], "same_as_first": sha == first,
                          "kernels_ms": {k: [c, round(t, 3)] for k, c, t in top}}), flush=True)
        res = None
    return 0


if __name__ == "__main__":
    sys.exit(main())

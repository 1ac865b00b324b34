#!/usr/bin/env python
"""bench.py — reads overlapped / s (+ alignment GCUPS) of the hash-overlap path on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload K1|K2|K3|K4|K5] [--scale f]
                  [--shape wide|default] [--impl hovl|reference]

One step = one pass of the whole path (splitter -> KmerTable -> pair counting -> dispatch ->
banded dovetail DP -> records) over the named synthetic read set.  `value` times it with the
reads already in HBM (hovl_run_device); `e2e` times the reference-facing call hovl_run on
pinned HOST buffers (H2D of the reads and D2H of the records inside the timed region).
`--impl reference` times the CPU restatement of the reference (oracle/, all host threads) on a
bounded sample of the same workload; the Scala/JVM original cannot run here (no JVM).
N > 1 (torchrun, one rank per GPU): see hovl/dist.py — k-mer-range sharded, two all-to-alls.
Every line carries `records_sha256` / `ovl_sha256` of the last timed step (device-resident and host-buffer
paths must agree, and the hashes must not depend on N), and the cpu_baseline leg doubles as a parity check:
the GPU runs the CPU leg's sample and its records must equal the oracle's (`parity`).
K1 / K2 are the reference's own read sets (c_ruddii rebuilt from the bank layout, crp177 with -k 16 -m HOXD1.txt)
in REFERENCE order with the 16-bit pair key, checked against the committed oracle fixtures.
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "sequence-aligner_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np

METRIC = "reads overlapped/sec"
UNIT = "reads/s"
# reads in the CPU leg's bounded sample (about 10-30 s of oracle time on 8 host threads)
CPU_SAMPLE_READS = {"K3": 24000, "K4": 200000, "K5": 400}
FIXTURES = ("K1", "K2")          # the reference's own read sets (tests/golden): reference order, 16-bit pair key
ALIGNER = "linear"
DIST_MODE = "partials"
DIST_DESC = {"partials": "reads sharded by id, k-mers by hash of the code, partial pair counts to the owner of fst",
             "replicate": "reads sharded by id, k-mer codes by range, sorted table gathered on every rank, pairs by owner of fst"}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="hovl", choices=["hovl", "reference"])
    ap.add_argument("--workload", default="K3", choices=["K1", "K2", "K3", "K4", "K5"])
    ap.add_argument("--scale", type=float, default=1.0, help="shrink genome and read count (tests)")
    ap.add_argument("--shape", default="wide", choices=["wide", "default"])
    ap.add_argument("--cpu-sample", type=float, default=0.0, help="fraction of the workload the CPU leg runs (0 = auto)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--opt", action="append", default=[], metavar="NAME=VALUE",
                    help="hovl_set_option tuning knob (results do not change), e.g. --opt gather_threads=64")
    ap.add_argument("--dist-mode", default="auto", choices=["auto", "partials", "replicate"],
                    help="N > 1: exchange partial pair counts (k-mer table sharded by hash), or gather the sorted table on every "
                         "rank and count the pairs of the rank's own reads (tables dominated by chance matches); auto = replicate for K4")
    ap.add_argument("--in-process", action="store_true",
                    help="--gpus N from ONE process: hovl_create(devices, N) + hovl_run (csrc/multi.cu, peer memory, no NCCL); "
                         "launch without torchrun.  Inputs are host buffers, so value == e2e")
    ap.add_argument("--aligner", default="linear", choices=["linear", "quadratic"],
                    help="linear = fast dovetail (Project4 default), quadratic = --quadratic-align")
    return ap.parse_args()


def load_fixture(name):
    """K1: c_ruddii, 32,000 x 100 bp reads rebuilt from the bank layout (tests/golden/make_golden.py), defaults (k=12).
    K2: crp177, 223 x 100 bp, -k 16 -m HOXD1.txt.  Returns (w, bases, offsets, settings kwargs, golden npz path)."""
    from hovl import io as hio
    G = os.path.join(ROOT, "tests", "golden")
    if name == "K1":
        lay = np.load(os.path.join(G, "c_ruddii_layout.npy")).astype(np.int64)
        contig, _ = hio.read_seq(os.path.join(G, "c_ruddii.fasta"))
        bases = np.ascontiguousarray(contig[lay[:, None] + np.arange(100)[None, :]].reshape(-1))
        offs = (np.arange(len(lay) + 1) * 100).astype(np.uint64)
        w = dict(name="K1", reads=len(lay), length=100, genome=len(contig), k=12, sub=0.0, fixture="c_ruddii")
        return w, bases, offs, {}, os.path.join(G, "c_ruddii_oracle_k12.npz")
    bases, offs = hio.read_seq(os.path.join(G, "crp177.seq"))
    w = dict(name="K2", reads=len(offs) - 1, length=100, genome=1209, k=16, sub=0.0, fixture="crp177")
    return w, bases, offs, {"k": 16, "cost": hio.read_hoxd(os.path.join(G, "HOXD1.txt"))}, os.path.join(G, "crp177_oracle_k16.npz")


def records_digest(records):
    """sha256 of the 52-byte records as returned, and of the .ovl text calcOverlaps would write."""
    import hovl
    recs = np.ascontiguousarray(records)
    return {"records_sha256": hashlib.sha256(recs.tobytes()).hexdigest(),
            "ovl_sha256": hashlib.sha256(hovl.format_ovl(recs)).hexdigest(), "n_records": int(len(recs))}


def workload_config(w, shape, n_gpus):
    if w.get("fixture"):
        return {"workload": "%s: the reference's %s reads, %d x %d bp" % (w["name"], w["fixture"], w["reads"], w["length"]),
                "flags": "Project4 defaults" if w["name"] == "K1" else "-k 16 -m HOXD1.txt",
                "pair_key": "reference (id_mode=0: (fst<<16)^snd, Trove iteration order)",
                "l2": "the whole working set fits the 126 MB L2; a 256 MB buffer is overwritten between steps" if w["name"] == "K2"
                      else "intermediate arrays exceed the L2 only in the pair stage; a 256 MB buffer is overwritten between steps",
                "parallelism": "1 GPU"}
    return {"workload": "%s: synthetic %d bp genome, %d x %d bp reads (%.0fx), %.0f%% substitutions, k=%d" % (
                w["name"], w["genome"], w["reads"], w["length"], w["reads"] * w["length"] / w["genome"],
                100 * w["sub"], w["k"]),
            "flags": ((("-k %d --max-ignore %d --max-collisions 2147483647" % (w["k"], w["length"])) if shape == "wide"
                       else "-k %d (Project4 defaults)" % w["k"]) + (" --quadratic-align" if ALIGNER == "quadratic" else "")),
            "pair_key": "wide (id_mode=1, canonical order)",
            "l2": "inputs and every intermediate array are larger than the 126 MB L2; no flush needed",
            "parallelism": "1 GPU" if n_gpus == 1 else "%d GPUs: %s" % (n_gpus, DIST_DESC.get(DIST_MODE, DIST_MODE))}


def load_alu_model():
    """profiles/alu_pipe.json: what the ALU-pipe roofline of the aligner is computed from, all of it measured on
    B200 and committed under profiles/ — the issue rate of the integer ALU pipe (tools/pipe_probe.cu: one warp
    instruction per `cycles_per_alu_inst` cycles per SM sub-partition) and, per DP kernel, the ALU-pipe
    instructions it executes per 32 cells (ncu sm__inst_executed_pipe_alu / (cells / 32)) with the capture named."""
    try:
        with open(os.path.join(ROOT, "profiles", "alu_pipe.json")) as f:
            return json.load(f)
    except (OSError, ValueError):
        return None


def alu_ceiling_gcups(sm_mhz, alu_inst_per_32cells, cycles_per_inst, sms=148):
    """cells/s one GPU can retire if its ALU pipes issue back to back: SMs x 4 sub-partitions x 32 lanes x clock
    / (ALU warp-instructions per 32 cells x cycles per instruction)."""
    return sms * 4 * 32 * sm_mhz * 1e6 / (alu_inst_per_32cells * cycles_per_inst) / 1e9


def dp_kernel_roofline(row, cells, sm_mhz, model):
    """ALU-pipe roofline of ONE aligner kernel (per GPU).  `frac` is against the kernel's own ALU-pipe instruction
    count (issue efficiency); `frac_bare_cell` against the fewest instructions the cell recurrence needs
    (distance from the best possible cell, not from this kernel's instruction stream)."""
    if not model or not cells or row["ms"] <= 0:
        return None
    name = row["kernel"].split(":", 1)[-1].strip("(").split("<")[0]
    k = model.get("kernels", {}).get(name)
    if not k:
        return None
    cyc = model["cycles_per_alu_inst"]
    achieved = cells / (row["ms"] * 1e6)                                    # GCUPS
    peak = alu_ceiling_gcups(sm_mhz, k["alu_inst_per_32cells"], cyc)
    bare = model.get("bare_cell", {}).get(k.get("cell", ""), None)
    out = {"kernel": row["kernel"], "bound": "int32-alu", "achieved": achieved / 1e3, "peak": peak / 1e3, "unit": "TCUPS",
           "frac": achieved / peak, "traffic": None, "alu_inst_per_32cells": k["alu_inst_per_32cells"],
           "cycles_per_alu_inst": cyc, "source": k.get("source"), "probe": model.get("probe"),
           "cells_per_step": cells, "sm_mhz": sm_mhz,
           "launches_per_step": row["launches"], "avg_launch_ms": row["ms"] / max(row["launches"], 1)}
    if bare:
        issue = 148 * 4 * 32 * sm_mhz * 1e6 / bare["inst_per_32cells"] / 1e9      # one warp instruction per cycle per sub-partition
        alu = alu_ceiling_gcups(sm_mhz, bare["alu_inst_per_32cells"], cyc)
        out["peak_bare_cell"] = min(issue, alu) / 1e3
        out["frac_bare_cell"] = achieved / min(issue, alu)
        out["bare_cell"] = bare
    return out


def dp_roofline(rows, gcups, sm_mhz, world, model, cells_by_kernel):
    """Whole aligner stage against the ALU-pipe ceiling, cells-weighted over its kernels (per GPU x world)."""
    if not model:
        return None
    t_min = 0.0                                   # time the stage would take with every kernel at its ALU ceiling
    for r in rows:
        f = dp_kernel_roofline(r, cells_by_kernel.get(r["kernel"]), sm_mhz, model)
        if f:
            t_min += f["cells_per_step"] / (f["peak"] * 1e12)
    cells = sum(v for v in cells_by_kernel.values() if v)
    if t_min <= 0 or not cells:
        return None
    peak = cells / t_min / 1e9 * world
    return {"bound": "int32 ALU pipe", "peak_gcups": peak, "frac": gcups / peak,
            "cycles_per_alu_inst": model["cycles_per_alu_inst"]}


# ------------------------------------------------------------------ clocks
class ClockSampler:
    """SM clock / power / throttle reasons sampled DURING the timed region: NVML in a thread of
    this process (nvidia-smi's own counters; a polling nvidia-smi child process holds driver
    locks long enough to stretch a 200 ms step by 10 %), nvidia-smi as the fallback."""
    BAD = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40, "sw_power_cap": 0x4}

    def __init__(self, index, period=0.1):
        self.index, self.period, self.rows, self.stop_flag, self.thread, self.err = index, period, [], False, None, None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.reasons_fn = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
                pynvml.nvmlDeviceGetCurrentClocksThrottleReasons
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception as e:                       # noqa: BLE001
            self.err = "nvml unavailable: %r" % (e,)

    def _pump(self):
        nv = self.nv
        while not self.stop_flag:
            try:
                self.rows.append((nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM), self.max_mhz, 0.0,
                                  self.reasons_fn(self.h)))
            except Exception as e:                   # noqa: BLE001
                self.err = repr(e)
                return
            time.sleep(self.period)

    def stop(self):
        self.stop_flag = True
        if self.thread:
            self.thread.join(timeout=2)
        if not self.rows:
            return self._smi_once()
        sm = [r[0] for r in self.rows]
        reasons = sorted(k for k, bit in self.BAD.items() if any(r[3] & bit for r in self.rows))
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(r[1] for r in self.rows)),
                "samples": len(sm), "reasons": reasons,
                "source": "nvml"}

    def _smi_once(self):
        try:
            out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=clocks.sm,clocks.max.sm",
                                  "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=10).stdout
            a, b = [float(x) for x in out.strip().split(",")[:2]]
            return {"sm_mhz": a, "sm_max_mhz": b, "samples": 1, "reasons": [], "source": "nvidia-smi after the run; " + str(self.err)}
        except Exception as e:                       # noqa: BLE001
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable: %r / %s" % (e, self.err)]}


# ------------------------------------------------------------------ CPU leg (oracle = restated reference)
def cpu_sample(args, shape):
    """(w, bases, offsets, oracle-settings kwargs, aligner) of the CPU leg's bounded sample: the same workload shrunk
    to `frac` of its genome and reads (same coverage, read length, error rate, k and flags); K1 / K2 run whole."""
    from hovl import synth
    if args.workload in FIXTURES:
        w, bases, offs, kw, _ = load_fixture(args.workload)
        return w, bases, offs, kw
    frac = args.cpu_sample
    if frac <= 0:
        frac = min(args.scale, CPU_SAMPLE_READS[args.workload] / synth.WORKLOADS[args.workload]["reads"])
    w = synth.scaled(args.workload, frac)
    bases, offs = synth.generate(w)
    return w, bases, offs, synth.shape_flags(w, shape)      # wide key + canonical order, like the full workload


def cpu_leg(args, shape, sample=None):
    """The reference's algorithm on the host cores: oracle/liboracle.so (C restatement of
    src/*.scala; kind "port" — the Scala original needs a JVM, absent here), threaded the way
    the JVM version is (parallel k-mer generation and block alignment, serial table and pair
    counting), on a bounded sample (cpu_sample)."""
    from oracle import oracle as O
    cores = os.cpu_count() or 1
    w, bases, offs, kw = sample or cpu_sample(args, shape)
    s = O.default_settings(**kw)
    s.aligner = 1 if ALIGNER == "quadratic" else 0
    t0 = time.perf_counter()
    r = O.run(s, bases, offs, threads=cores)
    dt = time.perf_counter() - t0
    if r.rc != 0:
        raise RuntimeError("oracle rc=%d" % r.rc)
    return {"value": w["reads"] / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "%s (%d reads, %d bp genome; same coverage/length/errors/k/flags), %.1f s, %d overlaps, %.3f GCUPS in align" % (
                w["name"], w["reads"], w["genome"], dt, r.stats["n_ovl_valid"],
                r.stats["dp_cells"] / max(r.ns["ns_align"], 1)),
            "seconds": dt, "reads": w["reads"], "stage_ns": r.ns, "records": r.records, "name": w["name"]}


def run_reference(args):
    from hovl import synth
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    w_full = load_fixture(args.workload)[0] if args.workload in FIXTURES else synth.scaled(args.workload, args.scale)
    sample = cpu_sample(args, args.shape)
    vals, last = [], None
    for i in range(args.warmup + args.steps):
        last = cpu_leg(args, args.shape, sample)
        if i >= args.warmup:
            vals.append(last)
    secs = sum(v["seconds"] for v in vals)
    value = sum(v["reads"] for v in vals) / secs
    cb = {k: last[k] for k in ("unit", "cores", "kind", "sample")}
    cb["value"] = value
    cfg = workload_config(w_full, args.shape, args.gpus)
    if last["name"] != w_full["name"]:                 # the line says what it ran: a bounded sample of the workload
        cfg["workload"] = "%s [each step = the bounded sample %s, %d reads]" % (cfg["workload"], last["name"], last["reads"])
    cfg["parallelism"] = "%d host threads (CPU arm)" % last["cores"]
    out = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / len(vals),
           "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int32",
           "data": "synthetic" if args.workload not in FIXTURES else "reference fixture", "config": cfg,
           "cpu_baseline": cb,
           "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    out.update(records_digest_oracle(last["records"]))
    print(json.dumps(out))


def records_digest_oracle(records):
    """Same digests as records_digest, for the CPU arm (formatted by the oracle's own writer)."""
    from oracle import oracle as O
    recs = np.ascontiguousarray(records)
    return {"records_sha256": hashlib.sha256(recs.tobytes()).hexdigest(),
            "ovl_sha256": hashlib.sha256(O.format_ovl(recs)).hexdigest(), "n_records": int(len(recs))}


# ------------------------------------------------------------------ roofline bookkeeping
def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def load_traffic(workload_name, kernel):
    """DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) of `kernel` on this workload from an
    `ncu` capture, kept in profiles/traffic.json as {"<workload>|<stage:kernel>": bytes}; None if never captured."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(path) as f:
            return json.load(f).get("%s|%s" % (workload_name, kernel))
    except (OSError, ValueError):
        return None


def kernel_rows(ktimes, st, key_bytes):
    """Per-kernel device time of ONE step with the algorithmic bytes each launch moves
    (DESIGN.md 'Kernels': per-unit figure x units per launch)."""
    nk, ninst, nb = st["n_kmers"], st["n_pair_instances"], st["n_bases"]
    tb = st.get("kmer_tuple_bytes") or 12                             # 12 B (code, read | loc) or 8 B (code, index)
    alg = {  # substring of "stage:kernel" -> (bound, algorithmic bytes per launch)
        "kmer_sort:(rs_pass_kernel": ("hbm", 2 * tb * nk),          # tuple read + written
        "kmer_sort:(rs_hist_kernel": ("hbm", 4 * nk),               # keys read once
        "pair_sort:(rs_pass_kernel": ("hbm", 2 * key_bytes * ninst),
        "pair_sort:(rs_hist_kernel": ("hbm", key_bytes * ninst),
        "expand:(expand_kernel": ("hbm", key_bytes * ninst),        # one key written per instance
        "encode:encode_kernel": ("hbm", nb // 4 + tb * nk),         # packed bases in, tuples out
        "pack:pack_kernel": ("hbm", nb + nb // 4),
        "pair_reduce:(scan_reduce_kernel": ("hbm", key_bytes * ninst),
        "pair_reduce:(scan_apply_kernel": ("hbm", key_bytes * ninst),
        "group:(scan_reduce_kernel": ("hbm", tb * nk),
        "group:(scan_apply_kernel": ("hbm", tb * nk + 8 * nk + 4 * nk),   # tuples in, section lists + bucket_of out
        "group:group_apply_kernel": ("hbm", tb * nk + int(9.6 * nk) + 8 * nk),   # tuples in, 1.2 list entries + (index, bucket) out
        "group:bucket_of_apply_kernel": ("hbm", 8 * nk + 4 * nk),
        # per-read gather: 8 B per section-list entry it has to look at (the kernel counts them: an instance is
        # read by the read that is fst and, unless the section thresholds already rule that out, by the
        # other one too), every k-mer looks up its bucket (4 B) and the six section bounds (24 B)
        "pair_gather:pair_gather_kernel": ("hbm", 8 * (st.get("n_list_entries") or 2 * ninst) + 28 * nk),
        "dp:(dp_kernel": ("int32", 0),
        "dp:(wb_kernel": ("int32", 0),
        "dp:(tp_phase1_kernel": ("int32", 0),
        "dp:(tp_phase2_kernel": ("int32", 0),
        "dp:(tp2_phase1_kernel": ("int32", 0),
        "dp:(tp2_phase2_kernel": ("int32", 0),
    }
    if not any(name.startswith("expand:(expand_kernel") for name in ktimes):
        # gather path: only the pairs that passed the count filter are sorted, (key, count) per pair
        alg["pair_sort:(rs_pass_kernel"] = ("hbm", 2 * (key_bytes + 4) * st["n_dispatched"])
        alg["pair_sort:(rs_hist_kernel"] = ("hbm", key_bytes * st["n_dispatched"])
    rows = []
    for name, (launches, ns) in ktimes.items():
        bound, b = None, None
        for pat, (bd, by) in alg.items():
            if name.startswith(pat):
                bound, b = bd, by
        if b and name.startswith("pair_gather:"):
            b = b / max(launches, 1)      # the figure above is for the whole read set; the launches (pilot + rest) split it
        rows.append({"kernel": name, "launches": launches, "ms": ns / 1e6, "bound": bound,
                     "alg_bytes_per_launch": b})
    rows.sort(key=lambda r: -r["ms"])
    return rows


# ------------------------------------------------------------------ GPU arm
def run_hovl(args):
    import torch
    import hovl
    from hovl import synth
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.in_process:
        return run_in_process(args)
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d (launch with torch.distributed.run)" % (args.gpus, world))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libhovl has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)

    fixture = args.workload in FIXTURES
    golden = None
    if fixture:
        if world != 1:
            raise SystemExit("K1 / K2 are single-GPU configs (reference order needs one table)")
        w, fb, fo, kw, golden = load_fixture(args.workload)
        n, L = w["reads"], w["length"]
        h_bases = torch.empty(len(fb), dtype=torch.uint8, pin_memory=True)
        h_bases.numpy()[:] = fb
        h_offs = torch.empty(n + 1, dtype=torch.int64, pin_memory=True)
        h_offs.numpy().view(np.uint64)[:] = fo
        bases, offs = h_bases.numpy(), h_offs.numpy().view(np.uint64)
        settings = hovl.default_settings(**kw)
    else:
        w = synth.scaled(args.workload, args.scale)
        n, L = w["reads"], w["length"]
        h_bases = torch.empty(n * L, dtype=torch.uint8, pin_memory=True)
        h_offs = torch.empty(n + 1, dtype=torch.int64, pin_memory=True)
        bases = h_bases.numpy()
        _, offs = synth.generate(w, out=bases)
        h_offs.numpy().view(np.uint64)[:] = offs
        offs = h_offs.numpy().view(np.uint64)
        settings = hovl.default_settings(**synth.shape_flags(w, args.shape))
    settings.aligner = 1 if args.aligner == "quadratic" else 0
    n_bases = int(offs[-1])

    stream = torch.cuda.Stream(device=dev)
    ctx = hovl.Context(local)
    ctx.set_option("stream", stream.cuda_stream)
    ctx.set_option("kernel_timing", 1)
    for kv in args.opt:
        name, value = kv.split("=", 1)
        ctx.set_option(name, int(value))

    if world == 1:
        d_bases = h_bases.to(dev)
        d_offs = h_offs.to(dev)
        torch.cuda.synchronize()

        def step_device():
            return ctx.run_device(settings, d_bases.data_ptr(), d_offs.data_ptr(), n, n_bases)

        def step_host():
            return ctx.run(settings, bases, offs)
    else:
        from hovl import dist as hdist
        dist_mode = args.dist_mode if args.dist_mode != "auto" else ("replicate" if args.workload == "K4" else "partials")
        eng = hdist.ShardedOverlapper(ctx, settings, stream, mode=dist_mode)
        lo, hi = hdist.read_slice(n, rank, world)
        d_bases = h_bases[lo * L:hi * L].to(dev)
        torch.cuda.synchronize()

        def step_device():
            return eng.run_device(d_bases, lo, hi, n, L)

        def step_host():
            return eng.run_host(h_bases[lo * L:hi * L], lo, hi, n, L)
    units_per_step = n

    def barrier():
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize()

    # K1 / K2 fit the 126 MB L2: a 256 MB buffer is overwritten between the timed steps (outside the timed
    # intervals: each step has its own event pair).  K3..K5 stream arrays larger than the L2; no flush.
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev) if fixture else None

    def timed(fn, k):
        barrier()
        ms, out = 0.0, None
        if flush is None:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(stream):
                e0.record(stream)
                for _ in range(k):
                    out = fn()
                e1.record(stream)
            barrier()
            ms = e0.elapsed_time(e1)
        else:
            for _ in range(k):
                with torch.cuda.stream(stream):
                    flush.add_(1)
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(stream)
                    out = fn()
                    e1.record(stream)
                barrier()
                ms += e0.elapsed_time(e1)
        if world > 1:
            import torch.distributed as dist
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, out

    prev = None
    for _ in range(args.warmup):
        prev = step_device()          # hold the previous result like the timed loop does: the library's
    del prev                          # page-locked result buffers reach their steady state during warm-up
    sampler = ClockSampler(local)
    sampler.start()
    ms_dev, res = timed(step_device, args.steps)
    clocks_dev = sampler.stop()
    ktimes = ctx.kernel_times() if world == 1 else res.ktimes
    st = res.stats
    st_local = st if world == 1 else res.local_stats          # rank 0's own counters, for its kernels' bytes
    digest = records_digest(res.records) if rank == 0 else None   # before the next run reuses the landing buffer
    res = None                                             # hand its result buffer back to the pool
    prev = step_host()                                     # warm-up of the host-buffer path
    prev = step_host()
    del prev
    sampler = ClockSampler(local)
    sampler.start()
    ms_e2e, res_h = timed(step_host, args.steps)
    clocks_e2e = sampler.stop()
    digest_e2e = records_digest(res_h.records) if rank == 0 else None
    n_rec_e2e = res_h.stats["n_records"]
    res_h = None

    # ---- parity leg (also the cpu_baseline): oracle and GPU on the same bounded sample, record for record
    cb, parity = None, None
    if not args.no_cpu:
        if rank == 0:
            sample = cpu_sample(args, args.shape)
            cb = cpu_leg(args, args.shape, sample)
            sw, sb, so, skw = sample
            s2 = hovl.default_settings(**skw)
            s2.aligner = settings.aligner
            got = ctx.run(s2, sb, so)
            ref = cb["records"]
            same = len(got.records) == len(ref) and np.ascontiguousarray(got.records).tobytes() == np.ascontiguousarray(ref).tobytes()
            parity = {"sample": cb["name"], "n_records": int(len(ref)),
                      "records_sha256": hashlib.sha256(np.ascontiguousarray(ref).tobytes()).hexdigest(),
                      "status": "ok: GPU records == CPU oracle records (all 13 ints, same order)" if same else "MISMATCH"}
            if golden:                                     # K1 / K2: the committed oracle vectors as well
                g = np.load(golden)["records"].view(hovl.RECORD_DTYPE).reshape(-1)
                parity["golden"] = os.path.relpath(golden, ROOT)
                parity["golden_status"] = "ok" if g.tobytes() == np.ascontiguousarray(got.records).tobytes() else "MISMATCH"
                same = same and parity["golden_status"] == "ok"
            got = None
            if not same:
                print(json.dumps({"error": "parity", "parity": parity}), file=sys.stderr)
        if world > 1:
            barrier()                                      # the other ranks wait for rank 0's CPU leg

    if rank != 0:
        return
    ms_step = ms_dev / args.steps
    value = units_per_step / (ms_step / 1e3)
    e2e_value = units_per_step / (ms_e2e / args.steps / 1e3)
    peak, peak_src = load_peaks()
    id_bits = max(1, int(n).bit_length())
    key_bytes = 4 if 2 * id_bits <= 32 else 8
    rows = kernel_rows(ktimes, st_local, key_bytes)          # rank 0's kernels against rank 0's units: per-GPU figures
    for r in rows:
        if r["bound"] == "hbm" and r["alg_bytes_per_launch"] and r["ms"] > 0:
            r["achieved_gbs"] = r["alg_bytes_per_launch"] * r["launches"] / (r["ms"] * 1e6)
            r["frac"] = r["achieved_gbs"] / peak
    sm_mhz = clocks_dev.get("sm_mhz") or 1965.0
    model = load_alu_model()
    # evaluated cells per aligner kernel (rank 0): phase-1 kernels share the evaluated phase-1 cells, the rest is phase 2
    p1_cells = st_local.get("dp_cells_phase1", 0) - st_local.get("dp_cells_skipped", 0)
    cells_eval_local = st_local["dp_cells"] - st_local.get("dp_cells_skipped", 0)
    cells_by_kernel = {}
    for r in rows:
        if r["bound"] != "int32":
            continue
        kname = r["kernel"]
        if "phase1" in kname:
            cells_by_kernel[kname] = p1_cells
        elif "phase2" in kname:
            cells_by_kernel[kname] = cells_eval_local - p1_cells
        else:
            cells_by_kernel[kname] = cells_eval_local       # lane-group / wide-band / quadratic: both phases in one kernel
    for r in rows:
        if r["bound"] == "int32":
            f = dp_kernel_roofline(r, cells_by_kernel.get(r["kernel"]), sm_mhz, model)
            if f:
                r["achieved_tcups"], r["frac"] = f["achieved"], f["frac"]
    top = rows[0] if rows else None

    def hbm_roofline(r):
        same_capture = world == 1 and not fixture
        return {"kernel": r["kernel"], "bound": "hbm", "achieved": r["achieved_gbs"], "peak": peak,
                "unit": "GB/s", "frac": r["frac"], "traffic": load_traffic(w["name"], r["kernel"]) if same_capture else None,
                "alg_bytes_per_launch": r["alg_bytes_per_launch"], "peak_source": peak_src,
                "launches_per_step": r["launches"], "avg_launch_ms": r["ms"] / max(r["launches"], 1),
                "share_of_step": r["ms"] / ms_step}
    top_hbm = next((r for r in rows if r["bound"] == "hbm" and "achieved_gbs" in r), None)
    roofline = None
    if top is not None and top["bound"] == "int32":           # the dominant kernel is an aligner kernel
        roofline = dp_kernel_roofline(top, cells_by_kernel.get(top["kernel"]), sm_mhz, model)
        if roofline:
            roofline["share_of_step"] = top["ms"] / ms_step
    if roofline is None and top_hbm:
        roofline = hbm_roofline(top_hbm)
    # dp_cells counts the cells of the reference's matrices; the pairs whose phase 1 was settled by the
    # exact-prefix probe did not evaluate theirs (dp_cells_skipped): GCUPS and the roofline use evaluated cells
    cells_eval = st["dp_cells"] - st.get("dp_cells_skipped", 0)
    gcups = cells_eval / max(st["ns_dp"], 1)
    gcups_ref = st["dp_cells"] / max(st["ns_dp"], 1)
    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
           "vs_baseline": None, "dtype": "int32", "data": "reference fixture" if fixture else "synthetic",
           "config": workload_config(w, args.shape, world),
           "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(n_bases + 8 * (n + 1)),
                   "d2h_bytes_per_step": int(52 * n_rec_e2e), "ms_per_step": ms_e2e / args.steps,
                   "records_sha256": digest_e2e["records_sha256"]},
           "records_sha256": digest["records_sha256"], "ovl_sha256": digest["ovl_sha256"],
           "records_match_e2e": digest["records_sha256"] == digest_e2e["records_sha256"],
           "gpu_launches": int(st["gpu_launches"]) * args.steps,
           "align_gcups": gcups,
           "dp": {"cells_per_step": cells_eval, "reference_cells_per_step": st["dp_cells"], "ms": st["ns_dp"] / 1e6,
                  "gcups": gcups, "gcups_reference_equivalent": gcups_ref,
                  "share_of_step": st["ns_dp"] / 1e6 / ms_step,
                  "roofline": dp_roofline(rows, cells_eval_local / max(st_local["ns_dp"], 1), sm_mhz, 1, model, cells_by_kernel)},
           "roofline": roofline,
           "roofline_hbm": hbm_roofline(top_hbm) if top_hbm else None,
           "kernels": rows[:12],
           "kernels_note": "rank 0's kernels with rank 0's units (per-GPU figures)" if world > 1 else "one GPU",
           "stage_ms": {k[3:]: st[k] / 1e6 for k in st if k.startswith("ns_")},
           "counts": {k: st[k] for k in ("n_reads", "n_kmers", "n_buckets", "max_bucket", "n_pair_instances",
                                         "n_distinct_pairs", "n_dispatched", "n_duds", "n_align_valid",
                                         "n_ovl_valid", "n_records")},
           "clocks": clocks_dev, "clocks_e2e": clocks_e2e}
    if top is not None:
        out["dominant_kernel"] = {"kernel": top["kernel"], "ms": top["ms"], "share_of_step": top["ms"] / ms_step}
    if cb is not None:
        out["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        out["parity"] = parity
    print(json.dumps(out))
    if parity is not None and not parity["status"].startswith("ok"):
        raise SystemExit("bench: GPU records differ from the CPU oracle's on the parity sample")


def run_in_process(args):
    """One process, one hovl_ctx over N devices (include/hovl.h: hovl_create with n_devices > 1): the call a JVM host makes.
    Reads sit in page-locked host memory; every step uploads the slices, shards the work through peer memory and lands
    the records in one page-locked buffer, so the timed figure is end to end by construction."""
    import torch
    import hovl
    from hovl import synth
    n_dev = args.gpus
    if torch.cuda.device_count() < n_dev:
        raise SystemExit("--in-process --gpus %d but only %d devices are visible" % (n_dev, torch.cuda.device_count()))
    w = synth.scaled(args.workload, args.scale)
    n, L = w["reads"], w["length"]
    h_bases = torch.empty(n * L, dtype=torch.uint8, pin_memory=True)
    bases = h_bases.numpy()
    _, offs = synth.generate(w, out=bases)
    settings = hovl.default_settings(**synth.shape_flags(w, args.shape))
    ctx = hovl.Context(list(range(n_dev)))
    for kv in args.opt:
        name, value = kv.split("=", 1)
        ctx.set_option(name, int(value))
    prev = None
    for _ in range(max(args.warmup, 1)):
        prev = ctx.run(settings, bases, offs)
    del prev
    sampler = ClockSampler(0)
    sampler.start()
    for d in range(n_dev):
        torch.cuda.synchronize(d)
    t0 = time.perf_counter()
    res = None
    for _ in range(args.steps):
        res = ctx.run(settings, bases, offs)           # blocking: returns when the records are in host memory
    dt = time.perf_counter() - t0
    clocks = sampler.stop()
    ms = 1e3 * dt / args.steps
    digest = records_digest(res.records)
    st = res.stats
    cells_eval = st["dp_cells"] - st.get("dp_cells_skipped", 0)
    cfg = workload_config(w, args.shape, n_dev)
    cfg["parallelism"] = "%d GPUs from ONE process: hovl_create(devices, %d) + hovl_run, exchanges through peer memory (csrc/multi.cu)" % (n_dev, n_dev)
    value = n / (ms / 1e3)
    print(json.dumps({"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_dev, "steps": args.steps, "warmup": args.warmup,
                      "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int32",
                      "data": "synthetic", "config": cfg, "timing": "host clock around blocking hovl_run calls (the call returns after its own device synchronisation)",
                      "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": int(n * L + 8 * (n + 1)),
                              "d2h_bytes_per_step": int(52 * st["n_records"]), "ms_per_step": ms},
                      "records_sha256": digest["records_sha256"], "ovl_sha256": digest["ovl_sha256"],
                      "gpu_launches": int(st["gpu_launches"]) * args.steps,
                      "align_gcups": cells_eval / max(st["ns_dp"], 1),
                      "stage_ms": {k[3:]: st[k] / 1e6 for k in st if k.startswith("ns_")},
                      "counts": {k: st[k] for k in ("n_reads", "n_kmers", "n_pair_instances", "n_distinct_pairs", "n_dispatched",
                                                    "n_duds", "n_align_valid", "n_ovl_valid", "n_records")},
                      "clocks": clocks}))


def _shutdown_dist():
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            dist.destroy_process_group()
    except Exception:
        pass


def main():
    global ALIGNER, DIST_MODE
    import atexit
    atexit.register(_shutdown_dist)
    args = parse_args()
    ALIGNER = args.aligner
    DIST_MODE = args.dist_mode if args.dist_mode != "auto" else ("replicate" if args.workload == "K4" else "partials")
    if args.impl == "reference":
        run_reference(args)
    else:
        run_hovl(args)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Prints the results table of BASELINE.md section 5 from the bench lines committed under profiles/.
  python profiles/baseline_table.py            # markdown on stdout
Every number in that table comes from one of the JSON lines named in its last column."""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))

# (row label, GPUs, file) — the line each row is read from
ROWS = [
    ("K1 c_ruddii 32,000 x 100 bp, k=12, defaults, reference order", 1, "r02_bench_K1_final.json"),
    ("K2 crp177 223 x 100 bp, k=16, -m HOXD1.txt, reference order", 1, "r02_bench_K2_final.json"),
    ("K3 5 Mbp / 200 k x 1 kb, k=16, wide flags", 1, "r02_bench_K3_final.json"),
    ("K3", 2, "r02_bench_K3_2gpu.json"),
    ("K3", 4, "r02_bench_K3_4gpu.json"),
    ("K3", 8, "r02_bench_K3_8gpu.json"),
    ("K3, one process (hovl_create n = 8)", 8, "r02_bench_K3_8gpu_inprocess.json"),
    ("K4 100 Mbp / 10 M x 150 bp, k=12, defaults (FULL size)", 1, "r02_bench_K4_final.json"),
    ("K4", 2, "r02_bench_K4_2gpu.json"),
    ("K4", 4, "r02_bench_K4_4gpu.json"),
    ("K4", 8, "r02_bench_K4_8gpu.json"),
    ("K5 x 0.01: 10 Mbp / 20 k x 10 kb, k=16, wide flags", 1, "r02_bench_K5x0.01_final.json"),
    ("K5 x 0.05", 4, "r02_bench_K5x0.05_4gpu.json"),
    ("K5 x 0.1: 100 Mbp / 200 k x 10 kb", 8, "r02_bench_K5x0.1_8gpu.json"),
]


def load(name):
    path = os.path.join(HERE, name)
    if not os.path.exists(path):
        return None
    for line in open(path):
        if line.startswith("{"):
            return json.loads(line)
    return None


def frac_of(d, pat):
    rows = [r for r in d.get("kernels", []) if pat in r["kernel"] and "frac" in r]
    if not rows:
        return "—"
    ms = sum(r["ms"] for r in rows)
    return "%.2f" % (sum(r["frac"] * r["ms"] for r in rows) / ms)


def main():
    print("| Config | GPUs | reads/s (HBM-resident) | e2e reads/s (host buffers in, records out) | ms / step | DP GCUPS | dominant kernel (share, fraction of its roofline) | sort / gather: fraction of HBM peak | CPU oracle reads/s (cores; sample) | records sha256 | parity | line |")
    print("|---|---|---|---|---|---|---|---|---|---|---|---|")
    for label, n, name in ROWS:
        d = load(name)
        if d is None:
            continue
        dom = d.get("dominant_kernel") or {}
        roof = d.get("roofline") or {}
        domtxt = "—"
        if dom:
            k = dom["kernel"].split(":", 1)[-1].strip("(").split("<")[0]
            domtxt = "%s (%.0f %%" % (k, 100 * dom["share_of_step"])
            if roof and roof.get("kernel") == dom["kernel"]:
                domtxt += ", %.2f of %s" % (roof["frac"], "ALU pipe" if roof["bound"] == "int32-alu" else "HBM")
                if "frac_bare_cell" in roof:
                    domtxt += "; %.2f of the bare-cell ceiling" % roof["frac_bare_cell"]
            domtxt += ")"
        cb = d.get("cpu_baseline")
        cbtxt = "—"
        if cb:
            cbtxt = "%s (%d; %s)" % ("{:,.0f}".format(cb["value"]), cb["cores"], cb["sample"].split(" (")[0])
        par = d.get("parity")
        partxt = "—"
        if par:
            partxt = ("GPU == oracle on %s (%s records)" % (par["sample"], "{:,}".format(par["n_records"]))) if par["status"].startswith("ok") else "MISMATCH"
            if par.get("golden_status") == "ok":
                partxt += "; == committed golden"
        print("| %s | %d | %s | %s | %.2f | %s | %s | %s / %s | %s | `%s` | %s | `%s` |" % (
            label, n, "{:,.0f}".format(d["value"]), "{:,.0f}".format(d["e2e"]["value"]), d["ms_per_step"],
            "{:,.0f}".format(d["align_gcups"]), domtxt, frac_of(d, "kmer_sort:(rs_pass"), frac_of(d, "pair_gather:"),
            cbtxt, d["records_sha256"][:12], partxt, name))


if __name__ == "__main__":
    sys.exit(main())

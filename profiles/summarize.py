#!/usr/bin/env python
"""Turn ncu outputs brought back in gpurun_out/ into the small text summaries committed here.
  python profiles/summarize.py launches gpurun_out/launches.csv > profiles/rNN_launches.txt
  python profiles/summarize.py full gpurun_out/prof_x.ncu-rep  > profiles/rNN_x_full.txt
"""
import collections
import csv
import io
import re
import subprocess
import sys

KEEP = re.compile(r"^(gpu__time_duration\.sum|dram__bytes_(read|write)\.sum|dram__throughput\.avg\.pct_of_peak_sustained_elapsed|"
                  r"gpu__dram_throughput\.avg\.pct_of_peak_sustained_elapsed|sm__throughput\.avg\.pct_of_peak_sustained_elapsed|"
                  r"launch__(registers_per_thread|grid_size|block_size|occupancy_limit_\w+|waves_per_multiprocessor)|"
                  r"sm__warps_active\.avg\.pct_of_peak_sustained_active|smsp__inst_executed\.sum|"
                  r"smsp__issue_active\.avg\.pct_of_peak_sustained_active|sm__inst_executed_pipe_(alu|fma|lsu|xu)\.avg\.pct_of_peak_sustained_active|"
                  r"sm__pipe_(alu|fma)_cycles_active\.avg\.pct_of_peak_sustained_active|l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum|"
                  r"sm__inst_executed\.sum|inst_executed|sm__cycles_active\.avg|sm__pipe_fmaheavy_cycles_active\.avg\.pct_of_peak_sustained_elapsed|smsp__sass_inst_executed_op_(global_ld|global_st|shared_ld|shared_st)\.sum|"
                  r"lts__t_sector_hit_rate\.pct|l1tex__t_sector_hit_rate\.pct|smsp__average_warps?_issue_stalled_\w+_per_issue_active\.ratio|"
                  r"smsp__warp_issue_stalled_\w+_per_warp_active\.pct|sm__cycles_elapsed\.max|smsp__cycles_active\.avg)$")


def launches(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    per = collections.OrderedDict()
    total = 0.0
    for r in csv.DictReader(lines):
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        name = re.sub(r"^void (hovl::)?", "", r["Kernel Name"])
        name = re.sub(r"\(.*$", "", name)
        ns = float(r["Metric Value"])
        d = per.setdefault(name, [0, 0.0])
        d[0] += 1
        d[1] += ns
        total += ns
    print("# kernel, launches, total ms, share of all profiled launches (cold-cache, serialised: compare SHARES)")
    for name, (n, ns) in sorted(per.items(), key=lambda kv: -kv[1][1]):
        print("%-70s %5d %10.3f ms %6.2f%%" % (name, n, ns / 1e6, 100 * ns / total))
    print("%-70s %5d %10.3f ms" % ("TOTAL", sum(v[0] for v in per.values()), total / 1e6))


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print("== %s  grid %s block %s" % (r[hdr.index("Kernel Name")], r[hdr.index("Grid Size")], r[hdr.index("Block Size")]))
        for h, u, v in zip(hdr, units, r):
            if KEEP.match(h):
                print("   %-90s %-14s %s" % (h, u, v))


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])

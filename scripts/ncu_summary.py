#!/usr/bin/env python
"""Summarise ncu outputs brought back in gpurun_out/ into small text files for profiles/.

  python scripts/ncu_summary.py launches <launches.csv>      per-kernel share of the step
  python scripts/ncu_summary.py raw <prof.ncu-rep> [regex]   selected raw metrics per captured launch
"""
import collections
import csv
import re
import subprocess
import sys

RAW_DEFAULT = (r"^(dram__bytes_(read|write)\.sum|gpu__time_duration\.sum|launch__registers_per_thread|"
               r"launch__grid_size|launch__block_size|sm__warps_active\.avg\.pct_of_peak_sustained_active|"
               r"sm__throughput\.avg\.pct_of_peak_sustained_elapsed|gpu__dram_throughput\.avg\.pct_of_peak_sustained_elapsed|"
               r"lts__t_bytes\.sum|l1tex__t_bytes\.sum|smsp__inst_executed\.sum|"
               r"sm__inst_executed_pipe_(fma|tensor[a-z_0-9]*)\.avg\.pct_of_peak_sustained_active|"
               r"sm__pipe_tensor[a-z_0-9]*cycles_active\.avg\.pct_of_peak_sustained_active|"
               r"smsp__average_warps_issue_stalled_(barrier|long_scoreboard|short_scoreboard|membar|wait|math_pipe_throttle|mio_throttle|not_selected)_per_issue_active\.ratio|"
               r"smsp__issue_active\.avg\.pct_of_peak_sustained_active)$")


def launches(path):
    rows = [r for r in csv.reader(open(path, errors="replace")) if len(r) > 10]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        v = float(r[vi].replace(",", ""))
        u = r[ui]
        ms = v / 1e6 if u.startswith("n") else (v / 1e3 if u.startswith("u") else (v * 1e3 if u in ("s", "second") else v))
        name = re.sub(r"\(.*", "", r[ki]).replace("<unnamed>::", "").replace("void ", "")[:70]
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += ms
    tot = sum(v[1] for v in agg.values())
    print("# %d launches, %.3f ms total device time (ncu-serialised, cold cache: compare SHARES)" % (sum(v[0] for v in agg.values()), tot))
    print("%-72s %5s %12s %8s %7s" % ("kernel", "n", "total ms", "avg ms", "share"))
    for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
        print("%-72s %5d %12.3f %8.3f %6.1f%%" % (k, v[0], v[1], v[1] / v[0], 100 * v[1] / tot))


def raw(path, pattern):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, u = rows[0], rows[1]
    rx = re.compile(pattern)
    for v in rows[2:]:
        d = dict(zip(h, v))
        print("## launch id %s  %s  grid %s block %s" % (d.get("ID"), re.sub(r"\(.*", "", d.get("Kernel Name", "")), d.get("Grid Size"), d.get("Block Size")))
        for name, unit, val in zip(h, u, v):
            if rx.search(name):
                print("%-100s %-14s %s" % (name, unit, val))


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2])
    else:
        raw(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else RAW_DEFAULT)

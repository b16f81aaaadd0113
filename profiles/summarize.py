#!/usr/bin/env python
"""Summarise an ncu launch list (gpu__time_duration.sum CSV) and, optionally, a --set full .ncu-rep.

    python profiles/summarize.py gpurun_out/launches.csv [gpurun_out/prof.ncu-rep ...] > profiles/rNN_x.txt
"""
import collections
import csv
import re
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum", "lts__t_bytes.sum", "l1tex__t_sector_hit_rate.pct",
        "lts__t_sector_hit_rate.pct", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__cycles_active.avg",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__m_xbar2l1tex_read_bytes.sum", "l1tex__m_l1tex2xbar_write_bytes.sum", "sm__cycles_elapsed.avg.per_second"]


def launches(path):
    rows = list(csv.DictReader(l for l in open(path) if l.startswith('"')))
    agg = collections.OrderedDict()
    for r in rows:
        k = re.sub(r"\(.*", "", r["Kernel Name"])[:70]
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += float(r["Metric Value"]) / 1e6
    tot = sum(v[1] for v in agg.values())
    print(f"# launch list {path}: {len(rows)} launches, {tot:.3f} ms total (cold-cache, serialised: compare shares)")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:70s} n={v[0]:4d} {v[1]:10.3f} ms {100 * v[1] / tot:5.1f}%  avg {1e3 * v[1] / v[0]:9.1f} us")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    print(f"\n# ncu --set full {path}")
    for r in rows[2:]:
        print(re.sub(r"\(.*", "", r[hdr.index("Kernel Name")])[:70])
        for w in WANT:
            if w in hdr:
                i = hdr.index(w)
                print(f"    {w:62s} {r[i]:>16s} {units[i]}")


if __name__ == "__main__":
    launches(sys.argv[1])
    for rep in sys.argv[2:]:
        full(rep)

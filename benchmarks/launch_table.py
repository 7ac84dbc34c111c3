#!/usr/bin/env python
"""Turns an ncu --csv launch list (gpu__time_duration.sum [+ dram__bytes_*]) into a per-launch table of the LAST step.

    python benchmarks/launch_table.py gpurun_out/launches.csv [--last N]"""
import collections
import csv
import sys


def main():
    path = sys.argv[1]
    last = int(sys.argv[sys.argv.index("--last") + 1]) if "--last" in sys.argv else 60
    lines = [l for l in open(path) if l.startswith('"')]
    d = collections.OrderedDict()
    for x in csv.DictReader(lines):
        e = d.setdefault(x["ID"], {"name": x["Kernel Name"], "grid": x["Grid Size"]})
        e[x["Metric Name"]] = float(x["Metric Value"].replace(",", ""))
    ids = list(d)[-last:]
    tot_t = tot_b = 0.0
    for i in ids:
        v = d[i]
        t = v.get("gpu__time_duration.sum", 0) / 1e3
        b = (v.get("dram__bytes_read.sum", 0) + v.get("dram__bytes_write.sum", 0)) / 1e6
        tot_t += t
        tot_b += b
        name = v["name"].replace("void ", "").replace("unnamed>::", "").replace("b2p::", "")
        print(f"{i:>4} {name[:58]:58} {v['grid']:>14} {t:8.1f} us {b:8.0f} MB")
    print(f"total {tot_t:.1f} us, {tot_b:.0f} MB DRAM over {len(ids)} launches")


if __name__ == "__main__":
    main()

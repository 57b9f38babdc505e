#!/usr/bin/env python
"""Refresh ONE family of profiles/traffic.json from a light ncu pass (long-format `--csv` of
`--metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum -k regex:<kernels>` over exactly one
eager bench step), when only that family's kernels changed since the full capture.

    python scripts/ncu_traffic_update.py METRICS.csv FAMILY CALLS_PER_STEP [workload]
"""
import collections
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    path, fam, calls = sys.argv[1], sys.argv[2], int(sys.argv[3])
    workload = sys.argv[4] if len(sys.argv) > 4 else "arxiv"
    by = collections.OrderedDict()
    for r in csv.reader(open(path)):
        if len(r) > 14 and r[0].isdigit():
            d = by.setdefault(r[0], {"name": r[4].split("(")[0]})
            d[r[12]] = float(r[14].replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3,
                                                         "us": 1.0, "ms": 1e3}.get(r[13], 1.0)
    ls = list(by.values())
    rd = sum(d["dram__bytes_read.sum"] for d in ls)
    wr = sum(d["dram__bytes_write.sum"] for d in ls)
    us = sum(d["gpu__time_duration.sum"] for d in ls)
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    traffic = json.load(open(tpath))
    traffic[workload][fam] = dict(
        bytes_per_launch=(rd + wr) / calls, read_bytes_per_call=rd / calls, write_bytes_per_call=wr / calls,
        ncu_us_per_call=us / calls, kernels_per_call=len(ls) / calls, calls_per_step=calls,
        source=f"profiles/{os.path.basename(path)} (ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,"
               f"gpu__time_duration.sum over the {len(ls)} {fam} launches of one eager bench step, after the kernels of "
               f"this family changed; the other families come from profiles/r02_ncu_step_raw.csv)")
    json.dump(traffic, open(tpath, "w"), indent=1)
    for d in ls:
        print(f"{d['gpu__time_duration.sum']:8.1f} us  {d['dram__bytes_read.sum'] / 1e6:8.1f} MB rd  {d['dram__bytes_write.sum'] / 1e6:8.1f} MB wr  {d['name'][:70]}")
    print(fam, (rd + wr) / calls / 1e6, "MB per call")


if __name__ == "__main__":
    main()

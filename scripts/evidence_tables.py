#!/usr/bin/env python
"""profiles/r02_ncu_agg_products.md and profiles/r02_ncu_vq_tensor_pipe.md from the CSVs scripts/gpu_r2_evidence.sh
brought back (gpurun_out/ncu_agg_products.csv, gpurun_out/ncu_vq_<K>_<D>.csv).  Needs no GPU."""
import collections
import csv
import glob
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
PEAK = 6543.7      # GB/s, MEASURED_PEAKS.json on this pool


def launches(path):
    by = collections.OrderedDict()
    for r in csv.reader(open(path)):
        if len(r) > 14 and r[0].isdigit():
            d = by.setdefault(r[0], {"name": r[4].split("(")[0], "grid": r[8]})
            d[r[12]] = float(r[14].replace(",", ""))
    return list(by.values())


def agg():
    ls = launches(os.path.join(OUT, "ncu_agg_products.csv"))
    # scripts/agg_bench.py runs 7 shapes, forward and backward, 7 timed repetitions each: keep the median launch of
    # every (kernel, grid) group in order of first appearance
    groups = collections.OrderedDict()
    for d in ls:
        groups.setdefault((d["name"], d["grid"]), []).append(d)
    shapes = {"(21168, 1, 1)": "arxiv-shaped 169,343 x 1,166,243, F = 128", "(2646, 1, 1)": "arxiv-shaped, F = 256",
              "(153065, 1, 1)": "products-shaped half graph 1,224,514 x 30,929,570, F = 100 (out: F = 100)",
              "(19134, 1, 1)": "products-shaped half graph, F = 256"}
    md = ["# Aggregation launches under ncu: measured DRAM bytes and L2 hit rate (round 2)", "",
          "`scripts/gpu_r2_evidence.sh`: `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,"
          "gpu__time_duration.sum,sm__warps_active... --clock-control none -k regex:aggregate_(kernel|stream_kernel) "
          "python scripts/agg_bench.py` on one B200; the median launch of every (kernel, grid) group "
          f"(forward and backward launches of a shape share a group when they run the same instantiation).  "
          f"Fraction = DRAM bytes / time / {PEAK:.0f} GB/s (measured copy peak).", "",
          "| kernel | grid | shape | launches | us | DRAM read MB | DRAM write MB | L2 hit % | warps active % | DRAM GB/s | of HBM peak |",
          "|---|---|---|---|---|---|---|---|---|---|---|"]
    for (name, grid), g in groups.items():
        g = sorted(g, key=lambda d: d["gpu__time_duration.sum"])
        d = g[len(g) // 2]
        us = d["gpu__time_duration.sum"] / 1e3
        rd, wr = d["dram__bytes_read.sum"] / 1e6, d["dram__bytes_write.sum"] / 1e6
        gbps = (rd + wr) / us * 1e3
        md.append(f"| `{name[:58]}` | {grid} | {shapes.get(grid, '')} | {len(g)} | {us:.1f} | {rd:.1f} | {wr:.1f} | "
                  f"{d['lts__t_sector_hit_rate.pct']:.0f} | {d['sm__warps_active.avg.pct_of_peak_sustained_active']:.0f} | "
                  f"{gbps:.0f} | {gbps / PEAK:.2f} |")
    open(os.path.join(ROOT, "profiles", "r02_ncu_agg_products.md"), "w").write("\n".join(md) + "\n")


def vq():
    md = ["# vq_tc_kernel (tcgen05 argmin-GEMM) under ncu: tensor-pipe activity, 1,000,000 rows (round 2)", "",
          "`scripts/gpu_r2_evidence.sh`: `ncu --metrics sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,"
          "gpu__time_duration.sum,dram__bytes_* --clock-control none -k regex:vq_tc_kernel -c 2 python scripts/vq_stats.py "
          "1000000 K D 1`; second launch of each point.  TFLOP/s = 2 * rows * K * D / kernel time; the whole "
          "assignment (search + gather + commitment partials) from the un-profiled run is beside it.", "",
          "| K | D | kernel us | tensor pipe active % | kernel TFLOP/s | of bf16 burst peak 1,665 | DRAM read MB | whole pass ms (no profiler) |",
          "|---|---|---|---|---|---|---|---|"]
    whole = {}
    pts = os.path.join(OUT, "vq_points.log")
    if os.path.exists(pts):
        for m in re.finditer(r"rows=(\d+) K=(\d+) D=(\d+) metric=\d+ path=2: ([\d.]+) ms", open(pts).read()):
            whole[(int(m.group(2)), int(m.group(3)))] = float(m.group(4))
    for path in sorted(glob.glob(os.path.join(OUT, "ncu_vq_*_*.csv")), key=lambda p: [int(x) for x in re.findall(r"\d+", os.path.basename(p))]):
        K, D = [int(x) for x in re.findall(r"\d+", os.path.basename(path))]
        ls = launches(path)
        if not ls:
            continue
        d = ls[-1]
        us = d["gpu__time_duration.sum"] / 1e3
        tf = 2 * 1e6 * K * D / (us * 1e-6) / 1e12
        md.append(f"| {K} | {D} | {us:.0f} | {d['sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active']:.1f} | {tf:.0f} | "
                  f"{tf / 1665.3:.2f} | {d['dram__bytes_read.sum'] / 1e6:.0f} | {whole.get((K, D), float('nan')):.3f} |")
    open(os.path.join(ROOT, "profiles", "r02_ncu_vq_tensor_pipe.md"), "w").write("\n".join(md) + "\n")


if __name__ == "__main__":
    agg()
    vq()

#!/usr/bin/env python
"""Turn the raw ncu CSV of one eager bench step into profiles/traffic.json (measured DRAM bytes per C-ABI call,
what bench.py's `roofline.traffic` / measured-bytes `frac` read) and a per-kernel table.

    python scripts/ncu_traffic.py RAW.csv SEQUENCE.json TAG [workload]

RAW.csv      : `ncu -i prof.ncu-rep --page raw --csv` of `ncu --set full -k regex:stemgnn ... bench.py --steps 1
               --no-graph` (scripts/gpu_r2_call2.sh); one row per kernel launch, in launch order
SEQUENCE.json: bench.py --dump-sequence of the same command: the step's C-ABI calls in order with the number of
               kernels each launched, which is how launches are attributed to call families
TAG          : profiles/<TAG>_ncu_step_kernels.md is written beside profiles/traffic.json
Needs no GPU."""
import collections
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLASS = [                       # call family -> regex its kernels' names must match
    ("csr_build", r"radix_|csr_|k0_"),
    ("aggregate", r"aggregate_"),
    ("row_scale", r"row_scale"),
    ("linear_wgrad", r"wgrad|colsum"),
    ("linear", r"lin_tc|lin_w_image"),
    ("bn_stats", r"col_reduce_kernel<0>|col_final_kernel<0>"),
    ("bn_act_forward", r"apply_kernel<(false|0)>"),
    ("bn_act_backward", r"col_reduce_kernel<1>|col_final_kernel<1>|apply_kernel<(true|1)>"),
    ("vq_assign", r"vq_(tc|gather|rescan|scan|sqerr|sqnorm)"),
    ("vq_code_stats", r"stats_|radix_|csr_rowptr|k0_"),
    ("vq_ema_apply", r"ema_"),
    ("vq_backward", r"vq_backward"),
]
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "%": 1.0}


def family(call):
    return "aggregate" if call.startswith("aggregate") else call


def main():
    raw, seq_path, tag = sys.argv[1:4]
    workload = sys.argv[4] if len(sys.argv) > 4 else "arxiv"
    rows = list(csv.reader(open(raw)))
    hdr, units, rows = rows[0], rows[1], [r for r in rows[2:] if r and r[0].isdigit()]
    col = {h: i for i, h in enumerate(hdr)}

    def val(r, name):
        i = col[name]
        try:
            return float(r[i]) * UNIT.get(units[i], 1.0)
        except ValueError:
            return float("nan")

    seq = json.load(open(seq_path))
    pattern = []
    for call, n, _ms in seq:
        pattern += [(family(call), dict(CLASS)[family(call)])] * int(n)
    names = [r[col["Kernel Name"]] for r in rows]
    start = None
    for p in range(0, len(names) - len(pattern) + 1):
        if all(re.search(rx, names[p + i]) for i, (_, rx) in enumerate(pattern)):
            start = p
            break
    if start is None:
        sys.exit("the launch list does not contain one whole step matching the call sequence")
    step = rows[start:start + len(pattern)]
    fam = collections.OrderedDict()
    table = []
    for (f, _), r in zip(pattern, step):
        rd, wr, us = val(r, "dram__bytes_read.sum"), val(r, "dram__bytes_write.sum"), val(r, "gpu__time_duration.sum")
        d = fam.setdefault(f, dict(read=0.0, write=0.0, us=0.0, launches=0))
        d["read"] += rd; d["write"] += wr; d["us"] += us; d["launches"] += 1
        table.append((f, r[col["Kernel Name"]].split("(")[0][:64], r[col["Grid Size"]], us, rd / 1e6, wr / 1e6,
                      val(r, "lts__t_sector_hit_rate.pct"), val(r, "sm__warps_active.avg.pct_of_peak_sustained_active"),
                      val(r, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
                      val(r, "sm__issue_active.avg.pct_of_peak_sustained_elapsed"),
                      val(r, "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio")))
    calls = collections.Counter(family(c) for c, _, _ in seq)
    src = (f"profiles/{tag}_ncu_step_raw.csv (ncu --set full --clock-control none of one eager bench step, "
           f"scripts/gpu_r2_call2.sh): dram__bytes_read.sum + dram__bytes_write.sum summed over the kernels of each "
           f"C-ABI call (attribution: bench.py --dump-sequence), divided by the calls per step")
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    traffic = json.load(open(tpath)) if os.path.exists(tpath) else {}
    out = {}
    for f, d in fam.items():
        out[f] = dict(bytes_per_launch=(d["read"] + d["write"]) / calls[f], read_bytes_per_call=d["read"] / calls[f],
                      write_bytes_per_call=d["write"] / calls[f], ncu_us_per_call=d["us"] / calls[f],
                      kernels_per_call=d["launches"] / calls[f], calls_per_step=calls[f], source=src)
    traffic[workload] = out
    json.dump(traffic, open(tpath, "w"), indent=1)
    md = [f"# ncu --set full, one eager bench step ({workload}-shaped), round-2 kernels", "",
          f"Source: `profiles/{tag}_ncu_step_raw.csv` rows {start}..{start + len(pattern) - 1} (one whole step located with "
          "the call sequence `bench.py --dump-sequence` wrote).  Times are ncu's serialised, cold-cache launch "
          "durations: the SHARE of the step is what compares with the live CUDA-event times in the bench line.", "",
          "## Per C-ABI call family", "",
          "| family | calls/step | kernels/call | ncu us/call | DRAM read MB/call | DRAM write MB/call | GB/s (DRAM bytes / ncu time) |",
          "|---|---|---|---|---|---|---|"]
    for f, d in fam.items():
        c = calls[f]
        md.append(f"| {f} | {c} | {d['launches'] / c:.1f} | {d['us'] / c:.1f} | {d['read'] / c / 1e6:.1f} | "
                  f"{d['write'] / c / 1e6:.1f} | {(d['read'] + d['write']) / d['us'] / 1e3:.0f} |")
    tot = sum(d["us"] for d in fam.values())
    md += ["", f"Sum of the library's kernels in the step under ncu: {tot:.0f} us.", "", "## Per kernel launch", "",
           "| family | kernel | grid | us | DRAM rd MB | DRAM wr MB | L2 hit % | warps active % | tensor pipe % | issue active % | long-scoreboard stall / issue |",
           "|---|---|---|---|---|---|---|---|---|---|---|"]
    for t in table:
        md.append("| {} | `{}` | {} | {:.1f} | {:.1f} | {:.1f} | {:.0f} | {:.0f} | {:.0f} | {:.0f} | {:.1f} |".format(*t))
    path = os.path.join(ROOT, "profiles", f"{tag}_ncu_step_kernels.md")
    open(path, "w").write("\n".join(md) + "\n")
    print(path)
    for f, d in out.items():
        print(f"{f:16s} {d['bytes_per_launch'] / 1e6:9.1f} MB/call  {d['ncu_us_per_call']:8.1f} us/call")


if __name__ == "__main__":
    main()

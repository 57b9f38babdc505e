#!/usr/bin/env python
"""profiles/r02_sass_tensor_kernels.md: the SASS lines that show which hardware paths the shipped library uses
(cuobjdump -sass of stem-gnn_b200/libstemgnn_b200.so, needs no GPU).  Per kernel: counts of the mnemonics the
profiling recipe names (UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UBLKCP = cp.async.bulk, UTMALDG = tensor-map TMA,
LDGSTS = cp.async, SYNCS = mbarrier) and the first occurrences with their addresses."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "stem-gnn_b200", "libstemgnn_b200.so")
PAT = re.compile(r"\b(UTCHMMA|UTCQMMA|UTCCP|LDTM|STTM|UBLKCP|UTMALDG|UTMASTG|LDGSTS|SYNCS|UTCBAR|HMMA|IMMA|ACQBULK|CCTL)\b[.\w]*")
WANT = ("vq_tc_kernel", "lin_tc_kernel", "wgrad_tc_kernel", "aggregate_hub_ring_kernel", "aggregate_stream_kernel",
        "aggregate_kernel")


def main():
    sass = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True, check=True).stdout
    out = ["# SASS evidence, round 2 (`python scripts/sass_excerpts.py`; cuobjdump -sass of the shipped .so, sm_100a)", ""]
    fn, per = None, collections.OrderedDict()
    for ln in sass.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            continue
        if fn is None or not any(w in fn for w in WANT):
            continue
        m = PAT.search(ln)
        if m:
            d = per.setdefault(fn, dict(counts=collections.Counter(), first={}))
            op = m.group(0)
            d["counts"][op] += 1
            d["first"].setdefault(op, ln.strip()[:150])
    seen_short = collections.Counter()
    for fn, d in per.items():
        short = fn.split("(")[0]
        fam = next(w for w in WANT if w in fn)
        # every tensor kernel; one representative per family for the (many) aggregation instantiations
        if fam.startswith("aggregate"):
            seen_short[fam] += 1
            if seen_short[fam] > 1:
                continue
        out.append(f"## `{short}`")
        out.append("")
        out.append("| mnemonic | count | first occurrence |")
        out.append("|---|---|---|")
        for op, n in sorted(d["counts"].items(), key=lambda kv: -kv[1]):
            out.append(f"| `{op}` | {n} | `{d['first'][op]}` |")
        out.append("")
    tot = collections.Counter()
    for d in per.values():
        for op, n in d["counts"].items():
            tot[op.split(".")[0]] += n
    out.append("## Totals over the listed kernel families (all instantiations)")
    out.append("")
    out.append(", ".join(f"{k} x{v}" for k, v in sorted(tot.items())))
    out.append("")
    out.append("No `UTMALDG`: operands staged by `cp.async.bulk` (1-D bulk copies of pre-swizzled tile images, "
               "`UBLKCP`) and by producer warps; no tensor map is built anywhere in the library.")
    path = os.path.join(ROOT, "profiles", "r02_sass_tensor_kernels.md")
    open(path, "w").write("\n".join(out) + "\n")
    print(path)


if __name__ == "__main__":
    sys.exit(main())

#!/usr/bin/env python3
"""Summarise an ncu launch list (`--metrics gpu__time_duration.sum --csv`) per kernel: launches, ms, share.
Usage: python tools/launch_summary.py gpurun_out/launches.csv > profiles/rNN_launches_summary.md"""
import collections
import csv
import re
import sys


def main():
    rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
    hdr = rows[0]
    ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        if len(r) <= iv or "gpu__time_duration" not in ",".join(r):
            continue
        ns = float(r[iv].replace(",", "")) * {"ns": 1.0, "us": 1e3, "ms": 1e6, "ns ": 1.0}.get(r[iu].strip(), 1.0)
        name = re.sub(r"\(.*", "", r[ik]).split("::")[-1]
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += ns
    tot = sum(v[1] for v in agg.values())
    print(f"total {tot / 1e6:.2f} ms over {sum(v[0] for v in agg.values())} launches\n")
    print("| kernel | launches | ms | share |\n|---|---|---|---|")
    for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{k}` | {n} | {ns / 1e6:.3f} | {100 * ns / tot:.1f} % |")


if __name__ == "__main__":
    main()

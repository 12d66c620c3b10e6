#!/usr/bin/env python3
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv --log-file X` launch list: first launches in order, then time
and launch count by kernel.   python tools/launch_summary.py gpurun_out/launches.csv profiles/rXX_launches_summary.txt "command" """
import csv
import io
import sys

src, out, cmd = sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else ""
txt = open(src).read()
rows = list(csv.DictReader(io.StringIO(txt[txt.index('"ID"'):])))
scale = {"ns": 1e-6, "nsecond": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "second": 1e3}
L = [(r["Kernel Name"], float(r["Metric Value"].replace(",", "")) * scale[r["Metric Unit"]]) for r in rows]
tot = sum(m for _, m in L)
o = [cmd, "(per-launch times under ncu are cold-cache and serialised: shares, not absolutes)", "", "first 14 launches in order:"]
o += [f"  {k:4d} {m:10.3f} ms  {n[:100]}" for k, (n, m) in enumerate(L[:14])]
o += ["", f"by kernel over the {len(L)} launches ({tot:.1f} ms):"]
agg = {}
for n, m in L:
    a = agg.setdefault(n, [0, 0.0])
    a[0] += 1
    a[1] += m
o += [f"  {c:5d} launches {m:10.3f} ms  {m / tot * 100:5.1f} %  {m / c * 1000:8.1f} us each  {n[:100]}"
      for n, (c, m) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:20]]
open(out, "w").write("\n".join(o) + "\n")
print("\n".join(o))

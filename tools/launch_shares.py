#!/usr/bin/env python
"""Per-kernel share of the summed launch durations in an ncu launch list (--metrics gpu__time_duration.sum --csv)."""
import csv
import sys
from collections import defaultdict

tot, cnt = defaultdict(float), defaultdict(int)
with open(sys.argv[1]) as f:
    rows = [r for r in csv.reader(f) if len(r) > 14 and r[0].isdigit()]
for r in rows:
    name = r[4].split("(")[0].replace("unnamed>::", "")
    tot[name] += float(r[14]) / 1e3
    cnt[name] += 1
s = sum(tot.values())
print(f"# {sys.argv[1]}: {len(rows)} launches, {s / 1e3:.3f} ms summed (cold-cache, serialised under ncu)")
for k in sorted(tot, key=tot.get, reverse=True):
    print(f"{k:28s} launches {cnt[k]:4d}  total {tot[k] / 1e3:9.3f} ms  share {tot[k] / s * 100:5.1f} %  avg {tot[k] / cnt[k]:8.1f} us")

#!/usr/bin/env python3
"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name: launches, total ms, share."""
import collections
import csv
import re
import sys

rows = list(csv.DictReader(l for l in open(sys.argv[1]) if l.startswith('"')))
tot = collections.Counter()
cnt = collections.Counter()
for r in rows:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = re.sub(r"\(.*", "", r["Kernel Name"])
    name = re.sub(r"<.*", "", name).split("::")[-1]
    v = float(r["Metric Value"].replace(",", ""))
    if r["Metric Unit"] in ("us", "usecond"):
        v *= 1e3
    elif r["Metric Unit"] in ("ms", "msecond"):
        v *= 1e6
    tot[name] += v
    cnt[name] += 1
allv = sum(tot.values())
for k, v in tot.most_common():
    print("%-44s %5d launches %10.3f ms  %5.1f%%" % (k, cnt[k], v / 1e6, 100 * v / allv))
print("%-44s %5d launches %10.3f ms" % ("total", sum(cnt.values()), allv / 1e6))

#!/usr/bin/env python3
"""Per-source-line hot spots per kernel of an .ncu-rep (needs -lineinfo + --import-source on).
Usage: ncu_lines.py report.ncu-rep [kernel-substring] [top=30]"""
import collections, csv, io, subprocess, sys
rep = sys.argv[1]; pat = sys.argv[2] if len(sys.argv) > 2 else ""; top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
cur = fun = hdr = None
agg = collections.OrderedDict()
done = set()
for r in csv.reader(io.StringIO(txt)):
    if r and r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r and r[0] == "Function Name":
        fun = r[1].split("(")[0]; continue
    if r and r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) < len(hdr) or r[0] == "" or pat not in (fun or ""): continue
    k = [hdr.index(x) for x in ("# Samples", "Instructions Executed", "Thread Instructions Executed")]
    try:
        v = (int(r[k[0]] or 0), int(r[k[1]] or 0), int(r[k[2]] or 0))
    except ValueError:
        continue
    d = agg.setdefault(fun, {})
    key = (cur, int(r[0]), r[1].strip()[:100])
    if key not in d: d[key] = v          # first captured launch of each kernel only
for fun, d in agg.items():
    ts = sum(v[0] for v in d.values()) or 1; ti = sum(v[1] for v in d.values()) or 1
    print("== %s: %d samples, %d warp instructions, %.1f threads/inst" % (fun, ts, ti, sum(v[2] for v in d.values()) / ti))
    for key, v in sorted(d.items(), key=lambda kv: -kv[1][1])[:top]:
        print("%5.1f%% smp %5.1f%% inst %5.1f thr  %s:%d  %s" % (100 * v[0] / ts, 100 * v[1] / ti, v[2] / max(1, v[1]), key[0], key[1], key[2]))
    b = collections.Counter(); bi = collections.Counter(); bt = collections.Counter()
    for (f, ln, src), v in d.items():
        kk = "%s:%d-%d" % (f, ln // 20 * 20, ln // 20 * 20 + 19)
        b[kk] += v[0]; bi[kk] += v[1]; bt[kk] += v[2]
    print("   -- 20-line bands")
    for kk, v in bi.most_common(14):
        print("%5.1f%% smp %5.1f%% inst %5.1f thr  %s" % (100 * b[kk] / ts, 100 * v / ti, bt[kk] / max(v, 1), kk))

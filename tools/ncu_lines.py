#!/usr/bin/env python3
"""Per-source-line hot spots of the first kernel in an .ncu-rep (needs -lineinfo + --import-source on).
Usage: ncu_lines.py report.ncu-rep [top=40]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
out, cur, hdr, nfun = [], None, None, 0
for r in rows:
    if r and r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r and r[0] == "Function Name":
        continue
    if r and r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) < len(hdr) or r[0] == "": continue
    k = [hdr.index(x) for x in ("# Samples", "Instructions Executed", "Thread Instructions Executed")]
    try:
        out.append((int(r[k[0]] or 0), int(r[k[1]] or 0), int(r[k[2]] or 0), cur, r[0], r[1].strip()[:110]))
    except ValueError:
        pass
ts = sum(o[0] for o in out) or 1; ti = sum(o[1] for o in out) or 1
print("total samples %d, warp instructions %d" % (ts, ti))
for s, i, t, f, ln, src in sorted(out, reverse=True)[:top]:
    print("%5.1f%% smp %5.1f%% inst %5.1f thr  %s:%s  %s" % (100 * s / ts, 100 * i / ti, t / max(i, 1), f, ln, src))
# ---- bucket view: file + 20-line bands
import collections
b = collections.Counter(); bi = collections.Counter(); bt = collections.Counter()
for s, i, t, f, ln, src in out:
    key = "%s:%d-%d" % (f, int(ln) // 20 * 20, int(ln) // 20 * 20 + 19)
    b[key] += s; bi[key] += i; bt[key] += t
print("---- 20-line bands")
for k, v in b.most_common(30):
    print("%5.1f%% smp %5.1f%% inst %5.1f thr  %s" % (100 * v / ts, 100 * bi[k] / ti, bt[k] / max(bi[k], 1), k))

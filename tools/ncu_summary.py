#!/usr/bin/env python3
"""Summarise an .ncu-rep (read here, no GPU needed) into a small text file for profiles/.
Usage: ncu_summary.py report.ncu-rep out.txt [kernel-regex]"""
import collections
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__inst_executed.sum",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio"]


def ncu(args):
    return subprocess.run(["ncu"] + args, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout


def main():
    rep, out = sys.argv[1], sys.argv[2]
    lines = []
    rows = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "raw", "--csv"]))))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        lines.append("== " + r[hdr.index("Kernel Name")])
        for k in KEYS:
            if k in hdr:
                lines.append("  %-85s %s %s" % (k, r[hdr.index(k)], units[hdr.index(k)]))
    rows = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "source", "--csv"]))))
    h = next((i for i, r in enumerate(rows) if "Source" in r and "Instructions Executed" in r), None)
    if h is not None:
        hd = rows[h]
        ie, te, src = hd.index("Instructions Executed"), hd.index("Thread Instructions Executed"), hd.index("Source")
        seen, cls, clst = set(), collections.Counter(), collections.Counter()
        for r in rows[h + 1:]:
            if len(r) <= te or r[0] in seen:      # first kernel only
                if len(r) > te and r[0] in seen:
                    break
                continue
            seen.add(r[0])
            try:
                n, t = int(r[ie] or 0), int(r[te] or 0)
            except ValueError:
                continue
            s = r[src].split()
            op = (s[1] if s and s[0].startswith("@") else s[0] if s else "?").split(".")[0]
            cls[op] += n
            clst[op] += t
        tot = sum(cls.values())
        lines.append("== SASS opcode mix of the first captured launch (warp instructions, avg active threads)")
        for op, c in cls.most_common(18):
            lines.append("  %-10s %5.1f%%   %.1f threads/inst" % (op, 100.0 * c / max(1, tot), clst[op] / max(1, c)))
    open(out, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines[:40]))


if __name__ == "__main__":
    main()

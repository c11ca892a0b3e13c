#!/usr/bin/env python3
"""profiles/r2_walk_counters.json and profiles/traffic.json from a full ncu capture of k_wave_trace_persistent at the bench
configuration (tools/gpu_profile.sh / tools/gpu_run*.sh write the .ncu-rep; this runs here, no GPU needed).
bench.py copies both into the roofline record (roofline.counters, roofline.traffic, roofline.hbm).
usage: make_walk_counters.py walk.ncu-rep "<how it was captured>" [summary file it is described in]"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
M = {
    "launch_ms": ("gpu__time_duration.sum", 1.0),
    "dram_bytes_read": ("dram__bytes_read.sum", 1.0),
    "dram_bytes_write": ("dram__bytes_write.sum", 1.0),
    "lanes_per_inst": ("smsp__thread_inst_executed_per_inst_executed.ratio", 1.0),
    "issue_active_pct": ("smsp__issue_active.avg.pct_of_peak_sustained_active", 1.0),
    "warps_active_pct": ("sm__warps_active.avg.pct_of_peak_sustained_active", 1.0),
    "l1_data_pipe_lsu_wavefronts_pct": ("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", 1.0),
    "l1_throughput_pct": ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
    "l1_hit_pct": ("l1tex__t_sector_hit_rate.pct", 1.0),
    "l2_hit_pct": ("lts__t_sector_hit_rate.pct", 1.0),
    "lts_throughput_pct": ("lts__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
    "dram_throughput_pct": ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
    "pipe_alu_pct": ("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", 1.0),
    "pipe_fma_pct": ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", 1.0),
    "stall_long_scoreboard_per_issue": ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", 1.0),
    "warp_inst": ("smsp__inst_executed.sum", 1.0),
    "sm_ghz": ("sm__cycles_elapsed.avg.per_second", 1.0),
}
UNIT = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1.0, "us": 1e-3, "s": 1e3, "ns": 1e-6,
        "Ghz": 1.0, "Mhz": 1e-3, "hz": 1e-9}


def main():
    rep, how = sys.argv[1], sys.argv[2]
    described = sys.argv[3] if len(sys.argv) > 3 else ""
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units, row = rows[0], rows[1], rows[2]
    out = {}
    for key, (name, _) in M.items():
        if name not in hdr:
            continue
        k = hdr.index(name)
        v = float(row[k].replace(",", ""))
        u = units[k]
        out[key] = v * UNIT.get(u, 1.0)
    out["dram_bytes_per_launch"] = out["dram_bytes_read"] + out["dram_bytes_write"]
    nsm, schedulers = 148, 4
    if "sm_ghz" in out and out["launch_ms"] > 0:
        ceiling = nsm * schedulers * out["sm_ghz"] * 1e9 * out["launch_ms"] * 1e-3       # warp instructions the chip could issue
        out["warp_inst_rate_frac_of_issue_ceiling"] = out["warp_inst"] / ceiling
        out["thread_inst_frac_of_ceiling"] = out["warp_inst"] * out["lanes_per_inst"] / (ceiling * 32.0)
    counters_path = os.path.join(ROOT, "profiles", "r2_walk_counters.json")
    old = json.load(open(counters_path)) if os.path.exists(counters_path) else {}
    new = {"kernel": "k_wave_trace_persistent", "how": how, "sorted": out}
    if "unsorted" in old:
        new["unsorted"] = old["unsorted"]
        new["unsorted_how"] = old.get("unsorted_how", old.get("how", ""))
    json.dump(new, open(counters_path, "w"), indent=1)
    traffic_path = os.path.join(ROOT, "profiles", "traffic.json")
    oldt = json.load(open(traffic_path)) if os.path.exists(traffic_path) else {}
    t = {"kernel": "k_wave_trace_persistent", "source": (described + " " if described else "") + "(" + how + ")",
         "dram_bytes_read": out["dram_bytes_read"], "dram_bytes_write": out["dram_bytes_write"],
         "dram_bytes_per_launch": out["dram_bytes_per_launch"], "launch_ms_under_ncu": out["launch_ms"]}
    for k in ("unsorted_queue", "note"):
        if k in oldt:
            t[k] = oldt[k]
    json.dump(t, open(traffic_path, "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Summarise an ncu report (read here, no GPU needed) into profiles/.

    python scripts/ncu_summary.py gpurun_out/<tag>_prof.ncu-rep profiles/<name>

Writes <name>_ncu_summary.json (key metrics per captured launch), <name>_ncu_hot.txt (the
hottest SASS instructions with their dominant stall reason).  The steady-state DRAM traffic that bench.py reports
(profiles/roofline_traffic.json) comes from scripts/ncu_steady.py, not from this cold, serialised capture.
"""
import csv
import io
import json
import os
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "launch__grid_size", "launch__block_size", "smsp__inst_executed.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
]
TO_BYTES = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def ncu(rep, *args):
    return subprocess.run(["ncu", "-i", rep, *args], check=True, capture_output=True, text=True).stdout


def main():
    rep, out = sys.argv[1], sys.argv[2]
    rows = list(csv.reader(io.StringIO(ncu(rep, "--page", "raw", "--csv"))))
    hdr, units, body = rows[0], rows[1], rows[2:]
    name_col = hdr.index("Kernel Name")
    launches = []
    for r in body:
        d = {"kernel": r[name_col]}
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                try:
                    d[k] = float(r[i])
                except ValueError:
                    d[k] = r[i]
                d[k + ".unit"] = units[i]
        launches.append(d)
    traffic = []
    for d in launches:
        if "dram__bytes_read.sum" in d:
            rd = d["dram__bytes_read.sum"] * TO_BYTES.get(d["dram__bytes_read.sum.unit"], 1)
            wr = d["dram__bytes_write.sum"] * TO_BYTES.get(d["dram__bytes_write.sum.unit"], 1)
            d["dram_bytes_total"] = rd + wr
            traffic.append(rd + wr)
    json.dump({"report": os.path.basename(rep), "launches": launches}, open(out + "_ncu_summary.json", "w"), indent=1)
    # one step of the mesher = an analyse + an expand launch: DRAM bytes of the first complete pair
    # (a pair whose expand kernel wrote nothing is the capacity probe of a fresh arena, not a step)
    pairs = []
    for a, b in zip(launches, launches[1:]):
        if "analyse" in a["kernel"] and "expand" in b["kernel"] and "dram_bytes_total" in a:
            if b["dram__bytes_write.sum"] * TO_BYTES.get(b["dram__bytes_write.sum.unit"], 1) > 1e6:
                pairs.append(a["dram_bytes_total"] + b["dram_bytes_total"])
    if pairs:
        traffic = pairs
    # (cold, serialised captures under-count DRAM traffic: profiles/roofline_traffic.json comes from scripts/ncu_steady.py)

    # hottest instructions of the first captured launch
    src = list(csv.reader(io.StringIO(ncu(rep, "--page", "source", "--csv", "--print-source", "sass"))))
    starts = [i for i, r in enumerate(src) if r and r[0] == "Address"]
    if starts:
        h = src[starts[0]]
        end = starts[1] - 1 if len(starts) > 1 else len(src)
        b = [r for r in src[starts[0] + 1:end] if len(r) == len(h)]
        si, ie = h.index("# Samples"), h.index("Instructions Executed")
        stall = [i for i, x in enumerate(h) if x.startswith("stall_") and "Not Issued" not in x]
        tot = sum(int(r[si]) for r in b if r[si].isdigit()) or 1
        with open(out + "_ncu_hot.txt", "w") as f:
            f.write(f"# {os.path.basename(rep)}: {len(b)} SASS instructions, {tot} warp samples, "
                    f"{sum(int(r[ie]) for r in b if r[ie].isdigit())} warp instructions executed\n")
            by = {}
            for r in b:
                for c in stall:
                    if r[c].isdigit():
                        by[h[c]] = by.get(h[c], 0) + int(r[c])
            f.write("# samples by stall reason: " + ", ".join(f"{k}={v}" for k, v in sorted(by.items(), key=lambda kv: -kv[1]) if v) + "\n")
            f.write("# idx  samples  %   executed  top stalls  | SASS\n")
            top = sorted(range(len(b)), key=lambda i: -int(b[i][si]) if b[i][si].isdigit() else 0)[:40]
            for i in sorted(top):
                r = b[i]
                st = sorted(((int(r[c]), h[c]) for c in stall if r[c].isdigit()), reverse=True)[:2]
                f.write(f"{i:5d} {r[si]:>7} {100 * int(r[si]) / tot:4.1f} {r[ie]:>9}  "
                        f"{st[0][1]}={st[0][0]},{st[1][1]}={st[1][0]} | {r[1].strip()}\n")
    print("wrote", out + "_ncu_summary.json")


if __name__ == "__main__":
    main()

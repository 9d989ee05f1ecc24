#!/usr/bin/env python
"""Turn the scratch ncu outputs under gpurun_out/ into the tracked summaries under profiles/.

    python profiles/summarise.py launches gpurun_out/X_launches.csv profiles/X_launches_summary.txt "<command>"
    python profiles/summarise.py full     gpurun_out/X_full.ncu-rep profiles/X_ncu_full.txt "<command>"
"""
import collections
import csv
import subprocess
import sys

FULL_METRICS = [
    "gpu__time_duration.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__grid_size",
    "launch__block_size", "lts__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
]


def launches(src, dst, cmd):
    rows = list(csv.reader(open(src)))
    h = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    ix = {k: j for j, k in enumerate(rows[h])}
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows[h + 1:]:
        if len(r) < len(rows[h]) or r[ix["Metric Name"]] != "gpu__time_duration.sum":
            continue
        v = float(r[ix["Metric Value"]].replace(",", ""))
        u = r[ix["Metric Unit"]]
        v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v * 1e6 if u == "s" else v
        k = r[ix["Kernel Name"]].split("(")[0][:90]
        agg[k][0] += 1
        agg[k][1] += v
    tot = sum(v[1] for v in agg.values())
    with open(dst, "w") as f:
        f.write(f"# command (under gpurun, 1x B200): {cmd}\n")
        f.write("# per-launch times are cold-cache and serialised under ncu: compare SHARES, not absolutes.\n")
        f.write(f"{'kernel':92s}{'launches':>9s}{'total_ms':>11s}{'share%':>8s}{'avg_us':>10s}\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:30]:
            f.write(f"{k:92s}{v[0]:9d}{v[1] / 1e3:11.3f}{100 * v[1] / tot:8.1f}{v[1] / v[0]:10.1f}\n")


def full(src, dst, cmd):
    raw = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    with open(dst, "w") as f:
        f.write(f"# command (under gpurun, 1x B200): {cmd}\n# one block per captured launch; from `ncu -i ... --page raw --csv`\n")
        for r in rows[2:]:
            f.write("----\n" + r[ix["Kernel Name"]][:120] + "\n")
            for m in FULL_METRICS:
                if m in ix:
                    f.write(f"  {m:86s}{r[ix[m]]:>16s} {units[ix[m]]}\n")


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](*sys.argv[2:5])

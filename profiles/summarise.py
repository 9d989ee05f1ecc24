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


# flops per lane of one executed FP32 SASS instruction (FMA = 2; packed ".F32x2" forms do two lanes' worth per thread)
FLOPS = {"FFMA2": 4, "FMUL2": 2, "FADD2": 2, "FFMA": 2, "FMUL": 1, "FADD": 1, "MUFU": 1, "FMNMX": 1}


def sass(src, dst, cmd):
    """Per-opcode EXECUTED instruction counts of the (first) kernel in an ncu report that was captured with
    --import-source on: the source page's per-SASS-line "# Instructions Executed" (warp level) and "# Thread
    Instructions Executed" columns, summed by opcode; the executed FP32 flop count follows from FLOPS above."""
    raw = subprocess.run(["ncu", "-i", src, "--page", "source", "--csv", "--print-source", "sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    h = next(i for i, r in enumerate(rows) if "Source" in r and any("Instructions Executed" in c for c in r))
    hdr = rows[h]
    i_src = hdr.index("Source")
    i_warp = next(j for j, c in enumerate(hdr) if c.strip() == "# Instructions Executed" or c.strip() == "Instructions Executed")
    i_thr = next((j for j, c in enumerate(hdr) if "Thread Instructions Executed" in c), None)
    agg = collections.defaultdict(lambda: [0, 0])
    for r in rows[h + 1:]:
        if len(r) <= max(i_src, i_warp):
            continue
        toks = r[i_src].replace("@", " @").split()
        toks = [t for t in toks if not t.startswith("@") and not t.startswith("!")]
        if not toks:
            continue
        op = toks[0].split(".")[0].rstrip(";")
        try:
            w = int(float(r[i_warp].replace(",", "") or 0))
            t = int(float(r[i_thr].replace(",", "") or 0)) if i_thr is not None else 32 * w
        except ValueError:
            continue
        agg[op][0] += w
        agg[op][1] += t
    tot_w = sum(v[0] for v in agg.values())
    flops = sum(FLOPS.get(op, 0) * v[1] for op, v in agg.items())
    with open(dst, "w") as f:
        f.write(f"# command (under gpurun, 1x B200): {cmd}\n")
        f.write("# executed SASS instructions of one captured launch by opcode (ncu source page); flops = thread-level "
                "count x flops per lane (FFMA2 4, FMUL2/FADD2 2, FFMA 2, FMUL/FADD/MUFU/FMNMX 1)\n")
        f.write(f"{'opcode':12s}{'warp_instr':>16s}{'share%':>8s}{'thread_instr':>18s}{'flops':>18s}\n")
        for op, v in sorted(agg.items(), key=lambda kv: -kv[1][0]):
            f.write(f"{op:12s}{v[0]:16d}{100.0 * v[0] / max(tot_w, 1):8.2f}{v[1]:18d}{FLOPS.get(op, 0) * v[1]:18d}\n")
        f.write(f"{'TOTAL':12s}{tot_w:16d}{100.0:8.2f}{sum(v[1] for v in agg.values()):18d}{flops:18d}\n")
    print(f"executed_flops_per_launch = {flops}")


if __name__ == "__main__":
    {"launches": launches, "full": full, "sass": sass}[sys.argv[1]](*sys.argv[2:5])

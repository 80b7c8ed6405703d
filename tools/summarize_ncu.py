#!/usr/bin/env python
"""Summarise ncu outputs into small tracked files under profiles/.

    python tools/summarize_ncu.py launches <launches.csv> <out.md> [title]
    python tools/summarize_ncu.py kernels  <raw.csv from `ncu -i x.ncu-rep --page raw --csv`> <out.md> [title]
"""
import collections
import csv
import re
import sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum",
        "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum", "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum"]


def short(name):
    name = re.sub(r"\(.*", "", name)
    return name.split("::")[-1]


def launches(path, out, title):
    lines = [l for l in open(path) if not l.startswith("==")]
    tot = collections.Counter(); cnt = collections.Counter()
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1e3 if u == "ns" else (v * 1e3 if u == "ms" else v)
        k = short(row["Kernel Name"]); tot[k] += v; cnt[k] += 1
    s = sum(tot.values())
    with open(out, "w") as f:
        f.write(f"# {title}\n\nncu `--metrics gpu__time_duration.sum --clock-control none` launch list "
                f"({sum(cnt.values())} launches, {s / 1e3:.2f} ms of kernel time; cold-cache, serialised: compare shares).\n\n")
        f.write("| kernel | launches | total us | avg us | share |\n|---|---:|---:|---:|---:|\n")
        for k, v in tot.most_common():
            f.write(f"| {k} | {cnt[k]} | {v:.1f} | {v / cnt[k]:.2f} | {100 * v / s:.1f}% |\n")


def kernels(path, out, title):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    with open(out, "w") as f:
        f.write(f"# {title}\n\n`ncu --set full --clock-control none --import-source on`, selected raw metrics per captured launch.\n\n")
        for r in rows[2:]:
            f.write(f"## {short(r[idx['Kernel Name']])}\n\n| metric | value | unit |\n|---|---:|---|\n")
            for k in KEYS:
                if k in idx:
                    f.write(f"| {k} | {r[idx[k]]} | {units[idx[k]]} |\n")
            f.write("\n")


if __name__ == "__main__":
    mode, src, dst = sys.argv[1:4]
    title = sys.argv[4] if len(sys.argv) > 4 else src
    (launches if mode == "launches" else kernels)(src, dst, title)

#!/usr/bin/env python
"""Turn ncu outputs into the small text summaries committed under profiles/.

    python profiles/summarize.py launches <launches.csv> > profiles/rNN_launches.md
    python profiles/summarize.py full <report.ncu-rep>   > profiles/rNN_full.md

`launches` = the `--metrics gpu__time_duration.sum` pass (per-kernel mean and share of the step;
cold-cache and serialised, so shares are what to read).  `full` = the raw page of a `--set full`
capture, reduced to the metrics DESIGN.md quotes.
"""
import collections
import csv
import subprocess
import sys

FULL = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__cycles_active.avg", "sm__cycles_elapsed.max",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_static", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "lts__t_sectors_op_red.sum", "lts__t_sectors_op_atom.sum"]


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        agg.setdefault(row["Kernel Name"], []).append(float(row["Metric Value"].replace(",", "")))
    ours = {k: v for k, v in agg.items() if "igs::" in k}
    step = sum(sum(v) / len(v) * (6 if "onesweep" in k else 1) for k, v in ours.items())
    print("| kernel | launches seen | mean us | share of the igs step |\n|---|---|---|---|")
    for k, v in agg.items():
        mean = sum(v) / len(v) / 1e3
        per_step = mean * (len(v) / max(1, min(len(x) for x in ours.values()))) if k in ours else None
        share = f"{per_step * 1e3 / sum(sum(x) / min(len(y) for y in ours.values()) for x in ours.values()):.1%}" \
            if k in ours else "(torch, outside the library)"
        print(f"| `{k[:90]}` | {len(v)} | {mean:.2f} | {share} |")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print(f"\n### {r[hdr.index('Kernel Name')]}  (launch id {r[0]})\n")
        for m in FULL:
            if m in hdr:
                i = hdr.index(m)
                print(f"- {m}: {r[i]} {units[i]}")


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])

#!/usr/bin/env python
"""Summarise an ncu report (read here, no GPU): launch metrics from `--page raw`, per-CUDA-line instruction and
stall-sample shares from `--page source --print-source cuda,sass`.

    python profiles/ncu_summarise.py gpurun_out/X.ncu-rep [--queries Q] [--top 30] [--kernel-index 0]
"""
import argparse
import csv
import io
import subprocess
import sys

RAW = ["gpu__time_duration.sum", "smsp__inst_executed.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
       "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
       "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
       "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "smsp__thread_inst_executed_per_inst_executed.ratio",
       "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
       "sm__inst_executed_pipe_fp64.sum", "smsp__inst_executed_pipe_fma.sum", "launch__waves_per_multiprocessor"]


def ncu(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True).stdout


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("--queries", type=float, default=0, help="divide instruction counts by this (e.g. queries per launch)")
    ap.add_argument("--top", type=int, default=30)
    ap.add_argument("--kernel", default="", help="substring of the kernel name whose source lines are listed (default: the first)")
    ap.add_argument("--instance", type=int, default=0, help="which launch of that kernel (0 = the first captured)")
    args = ap.parse_args()
    rows = list(csv.reader(io.StringIO(ncu(["-i", args.report, "--page", "raw", "--csv"]))))
    hdr, units = rows[0], rows[1]
    print("## launches\n")
    for r in rows[2:]:
        print("*", r[hdr.index("Kernel Name")][:90])
        for w in RAW:
            if w in hdr:
                i = hdr.index(w)
                print(f"    {w} = {r[i]} {units[i]}")
    src = list(csv.reader(io.StringIO(ncu(["-i", args.report, "--page", "source", "--csv", "--print-source", "cuda,sass"]))))
    # the export repeats per kernel instance; take the first function block
    lines, cur, seen_fn = {}, None, 0
    total_inst = total_samp = 0
    col = None
    active = False
    seen = -1
    cur_file = ""
    for r in src:
        if len(r) >= 2 and r[0] in ("File Path", "File Name"):
            cur_file = r[1]
            continue
        if len(r) >= 2 and r[0] == "Function Name":
            if active:
                break
            main = cur_file.endswith("_kernels.cu")          # the export has one block per (launch, source file)
            if args.kernel in r[1] and main:
                seen += 1
            active = args.kernel in r[1] and main and seen == args.instance
            col = None
            continue
        if not active:
            continue
        if len(r) > 8 and r[0] == "Line No":
            col = {name: i for i, name in enumerate(r)}
            continue
        if col is None or len(r) < 8:
            continue
        if r[0] not in ("", "-"):
            try:
                cur = int(r[0])
            except ValueError:
                continue
            if "Instructions Executed" not in col or "# Samples" not in col:
                continue
            inst = int(r[col["Instructions Executed"]] or 0)
            samp = int(r[col["# Samples"]] or 0)
            thr = int(r[col["Thread Instructions Executed"]] or 0)
            e = lines.setdefault(cur, {"src": r[1].strip(), "inst": 0, "samp": 0, "thr": 0})
            e["inst"] += inst; e["samp"] += samp; e["thr"] += thr
            total_inst += inst; total_samp += samp
    q = args.queries or 1.0
    print(f"\n## source lines (first kernel instance): {total_inst} warp instructions"
          + (f" = {total_inst / q:.1f} per query" if args.queries else "") + f", {total_samp} stall samples\n")
    print("    line | instr" + ("/query" if args.queries else "") + " | % instr | % samples | lanes | source")
    for ln, e in sorted(lines.items(), key=lambda kv: -kv[1]["samp"])[: args.top]:
        lanes = e["thr"] / e["inst"] if e["inst"] else 0
        print(f"    {ln:5d} | {e['inst'] / q:9.1f} | {100 * e['inst'] / max(total_inst, 1):5.1f} | {100 * e['samp'] / max(total_samp, 1):5.1f} | {lanes:4.1f} | `{e['src'][:110]}`")


if __name__ == "__main__":
    main()

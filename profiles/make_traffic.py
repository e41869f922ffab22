#!/usr/bin/env python
"""profiles/traffic.json from ncu reports: DRAM bytes (read + write) and duration per launch of every captured kernel,
keyed by (kernel, workload, sha of the CUDA sources the captured binary was built from).  bench.py reads it for
`roofline.traffic` and prints null + `traffic_stale: true` when the sha differs from the sources it runs.

    python profiles/make_traffic.py c3=gpurun_out/r02v_c3_rows.ncu-rep c3=gpurun_out/r02v_c3_aux.ncu-rep c5=...
"""
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def rows_of(report):
    out = subprocess.run(["ncu", "-i", report, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        yield {h: (r[i], units[i]) for i, h in enumerate(hdr)}


def to_bytes(val, unit):
    v = float(val.replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]


def main():
    import bench
    shas = {}
    entries = {}
    for arg in sys.argv[1:]:
        workload, report = arg.split("=", 1)
        for r in rows_of(report):
            name = re.sub(r"\(.*", "", r["Kernel Name"][0]).replace("void ", "").replace("mmb::", "")
            base = re.sub(r"<.*", "", name)
            rd = to_bytes(*r["dram__bytes_read.sum"]); wr = to_bytes(*r["dram__bytes_write.sum"])
            dur = float(r["gpu__time_duration.sum"][0].replace(",", ""))
            dur_us = dur if r["gpu__time_duration.sum"][1] == "us" else dur * {"ns": 1e-3, "ms": 1e3}[r["gpu__time_duration.sum"][1]]
            if base not in shas:
                shas[base] = bench.kernel_source_sha(base)
            e = entries.setdefault((base, workload), {"kernel": base, "instance": name, "workload": workload, "source_sha": shas[base],
                                                      "source_file": bench.kernel_source_file(base),
                                                      "capture": "profiles/" + os.path.basename(report).replace(".ncu-rep", "_summary.md"),
                                                      "launches": 0, "dram_read": 0.0, "dram_write": 0.0, "duration_us": 0.0,
                                                      "inst": 0.0})
            e["launches"] += 1; e["dram_read"] += rd; e["dram_write"] += wr; e["duration_us"] += dur_us
            e["inst"] += float(r["smsp__inst_executed.sum"][0].replace(",", ""))
    out = []
    for e in entries.values():
        n = e.pop("launches")
        e["launches_captured"] = n
        e["dram_bytes_per_launch"] = (e["dram_read"] + e["dram_write"]) / n
        e["dram_read_per_launch"] = e.pop("dram_read") / n
        e["dram_write_per_launch"] = e.pop("dram_write") / n
        e["duration_us_under_ncu"] = e.pop("duration_us") / n
        e["warp_instructions_per_launch"] = e.pop("inst") / n
        out.append(e)
    path = os.path.join(ROOT, "profiles", "traffic.json")
    old = []
    if os.path.exists(path):
        old = [e for e in json.load(open(path)).get("entries", [])
               if (e["kernel"], e["workload"]) not in {(x["kernel"], x["workload"]) for x in out}]
    json.dump({"how": "ncu --set full --clock-control none (cold cache, serialised), one or more launches per kernel; "
                      "profiles/make_traffic.py", "entries": old + out}, open(path, "w"), indent=1)
    for e in out:
        print(f"{e['kernel']:28s} {e['workload']}  {e['dram_bytes_per_launch'] / 1e6:8.2f} MB  {e['duration_us_under_ncu']:8.1f} us  "
              f"{e['warp_instructions_per_launch'] / 1e6:7.2f} M instr")


if __name__ == "__main__":
    main()

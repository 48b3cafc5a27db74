#!/usr/bin/env python3
"""Summarises an ncu launch list (--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv) into
per-kernel totals (launches, time, share of the captured time, DRAM bytes per launch) and writes profiles/traffic.json
(the measured DRAM bytes per launch of the extend kernel, which bench.py reports as roofline.traffic).

    python tools/summarise_launches.py profiles/r01g_launches_bench_c2.csv [--tag r01g]
"""
import argparse
import csv
import json
import os
import re
import sys
from collections import defaultdict

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("--tag", default="")
    ap.add_argument("--command", default="ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 1200 --csv python bench.py --steps 2 --warmup 1")
    a = ap.parse_args()
    rows = []
    with open(a.csv) as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        rows.append(r)
    per = defaultdict(lambda: defaultdict(float))
    ids = defaultdict(set)
    for r in rows:
        name = re.sub(r"\(.*", "", r["Kernel Name"])
        v = float(r["Metric Value"].replace(",", ""))
        m = r["Metric Name"]
        if m == "gpu__time_duration.sum":
            unit = r["Metric Unit"]
            v *= {"ns": 1.0, "us": 1e3, "usecond": 1e3, "ms": 1e6, "msecond": 1e6, "nsecond": 1.0, "second": 1e9}.get(unit, 1.0)
        if m.startswith("dram__bytes"):
            unit = r["Metric Unit"]
            v *= {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1.0)
        per[name][m] += v
        ids[name].add(r["ID"])
    total_ns = sum(p["gpu__time_duration.sum"] for p in per.values())
    table = []
    for name, p in sorted(per.items(), key=lambda kv: -kv[1]["gpu__time_duration.sum"]):
        n = len(ids[name])
        table.append(dict(kernel=name, launches=n, total_ms=p["gpu__time_duration.sum"] / 1e6, share=p["gpu__time_duration.sum"] / total_ns,
                          mean_us=p["gpu__time_duration.sum"] / n / 1e3, dram_read_per_launch=p["dram__bytes_read.sum"] / n,
                          dram_write_per_launch=p["dram__bytes_write.sum"] / n))
    out = dict(source=os.path.relpath(a.csv, ROOT), command=a.command, captured_launches=sum(len(v) for v in ids.values()),
               captured_ms=total_ns / 1e6, kernels=table)
    base = os.path.splitext(a.csv)[0] + "_summary.json"
    json.dump(out, open(base, "w"), indent=1)
    print("wrote", base)
    for t in table[:12]:
        print("%-40s n=%5d  %8.3f ms  %5.1f %%  %9.1f us/launch  DRAM %.1f MB/launch" % (t["kernel"][:40], t["launches"], t["total_ms"], 100 * t["share"],
                                                                                     t["mean_us"], (t["dram_read_per_launch"] + t["dram_write_per_launch"]) / 1e6))
    tr = [t for t in table if "trace_closest_persistent_kernel" in t["kernel"]]
    if tr:
        t = tr[0]
        traffic = dict(trace_closest_bytes_per_launch=t["dram_read_per_launch"] + t["dram_write_per_launch"],
                       dram_read_bytes_per_launch=t["dram_read_per_launch"], dram_write_bytes_per_launch=t["dram_write_per_launch"],
                       kernel="trace_closest_persistent_kernel", launches_averaged=t["launches"], mean_launch_ns_under_ncu=t["mean_us"] * 1e3,
                       share_of_captured_time=t["share"], command=a.command, source=os.path.relpath(a.csv, ROOT),
                       note="DRAM traffic of the extend stage is the ray batch in and the hit records out (48 B per ray of the launch); "
                            "the scene stays in L2/L1, so DRAM bytes are a few per cent of the algorithmic traversal bytes.")
        json.dump(traffic, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
        try:
            print("wrote profiles/traffic.json:", round(traffic["trace_closest_bytes_per_launch"] / 1e6, 1), "MB per launch, share", round(t["share"], 3))
        except BrokenPipeError:
            pass


if __name__ == "__main__":
    sys.exit(main())

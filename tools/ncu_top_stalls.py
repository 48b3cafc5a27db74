#!/usr/bin/env python3
"""Top stall locations of the captured kernels: python tools/ncu_top_stalls.py report.ncu-rep [N] [kernel-substring]
(reads ncu's source page; one table per captured launch)."""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 20
want = sys.argv[3] if len(sys.argv) > 3 else ""
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
heads = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
for n, hi in enumerate(heads):
    end = heads[n + 1] - 1 if n + 1 < len(heads) else len(rows)
    hdr = rows[hi]
    name = rows[hi - 1][1] if hi > 0 and len(rows[hi - 1]) > 1 else "?"
    if want and want not in name:
        continue
    ix = {h: i for i, h in enumerate(hdr)}
    data = [r for r in rows[hi + 1:end] if len(r) == len(hdr)]
    tot = sum(int(r[ix["# Samples"]] or 0) for r in data)
    print("kernel:", name[:100])
    print("total samples", tot, "instructions", len(data))
    reasons = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    agg = {h: sum(int(r[ix[h]] or 0) for r in data) for h in reasons}
    print("by reason:", ", ".join("%s %.1f%%" % (k[6:], 100.0 * v / max(tot, 1)) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
    for r in sorted(data, key=lambda r: -int(r[ix["# Samples"]] or 0))[:top_n]:
        c = int(r[ix["# Samples"]] or 0)
        st = sorted(((k[6:], int(r[ix[k]] or 0)) for k in reasons), key=lambda kv: -kv[1])[:2]
        print("%s %5.1f%%  %-64s %s" % (r[0][-5:], 100.0 * c / max(tot, 1), r[ix["Source"]][:64], st))
    print()

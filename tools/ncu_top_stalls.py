#!/usr/bin/env python3
"""Top stall locations of one captured kernel: python tools/ncu_top_stalls.py report.ncu-rep [N]  (reads ncu's source page)."""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 20
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]
ix = {h: i for i, h in enumerate(hdr)}
data = rows[hi + 1:]
tot = sum(int(r[ix["# Samples"]] or 0) for r in data)
print("kernel:", rows[0][1][:100] if rows[0] else "?")
print("total samples", tot, "instructions", len(data))
reasons = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = {h: sum(int(r[ix[h]] or 0) for r in data) for h in reasons}
print("by reason:", ", ".join("%s %.1f%%" % (k[6:], 100.0 * v / max(tot, 1)) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
for r in sorted(data, key=lambda r: -int(r[ix["# Samples"]] or 0))[:top_n]:
    n = int(r[ix["# Samples"]] or 0)
    st = sorted(((k[6:], int(r[ix[k]] or 0)) for k in reasons), key=lambda kv: -kv[1])[:2]
    print("%s %5.1f%%  %-64s %s" % (r[0][-5:], 100.0 * n / max(tot, 1), r[ix["Source"]][:64], st))

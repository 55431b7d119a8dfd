#!/usr/bin/env python
"""Top stalled instructions of the first kernel in an .ncu-rep captured with --import-source on.
Usage: python tools/ncu_top_stalls.py <report.ncu-rep> [count] > profiles/<name>_top_stalls.md"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
count = int(sys.argv[2]) if len(sys.argv) > 2 else 20
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
name = rows[0][1] if rows and len(rows[0]) > 1 else "?"
hdr = rows[1]
data, prev = [], -1
for r in rows[2:]:
    if len(r) != len(hdr) or not r[0].startswith("0x"):
        continue
    a = int(r[0], 16)
    if a < prev:
        break                                   # second launch in the report
    data.append(r)
    prev = a
total = sum(int(r[2]) for r in data)
first_reason = hdr.index("stall_barrier") if "stall_barrier" in hdr else 37
print(f"# top stalled instructions of `{name[:90]}`\n\n{rep.split('/')[-1]}: {len(data)} SASS instructions, {total} warp-stall samples.\n")
print("| # | samples | share | instruction | main stall reasons |\n|---|---|---|---|---|")
for r in sorted(data, key=lambda r: -int(r[2]))[:count]:
    reasons = sorted(((int(r[i]), hdr[i]) for i in range(first_reason, len(hdr)) if r[i].isdigit() and int(r[i]) > 0 and "Not Issued" not in hdr[i]), reverse=True)[:2]
    print(f"| {data.index(r)} | {r[2]} | {100 * int(r[2]) / total:.1f} % | `{r[1].strip()}` | {', '.join(f'{h} {c}' for c, h in reasons)} |")

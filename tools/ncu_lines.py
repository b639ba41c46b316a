"""Aggregate an `ncu --page source --csv --print-source cuda,sass` dump per CUDA source line.
Usage: ncu_lines.py dump.csv [top]"""
import csv
import sys
from collections import defaultdict

top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
agg = defaultdict(lambda: [0, 0, ""])
fname = ""
hdr = None
for r in csv.reader(open(sys.argv[1])):
  if not r:
    continue
  if r[0] == "File Path":
    fname = r[1].split("/")[-1]; continue
  if r[0] == "Function Name":
    continue
  if r[0] == "Line No":
    hdr = r; si = hdr.index("# Samples"); ii = hdr.index("Instructions Executed"); continue
  if hdr is None or r[2] != "-":     # keep only the per-source-line summary rows (Address == "-")
    continue
  try:
    s = int(r[si]); n = int(r[ii])
  except ValueError:
    continue
  k = (fname, int(r[0]))
  agg[k][0] += s; agg[k][1] += n; agg[k][2] = r[1].strip()[:110]
tot = sum(v[0] for v in agg.values()) or 1
toti = sum(v[1] for v in agg.values()) or 1
print(f"total samples {tot}  total warp-instructions {toti}")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
  print(f"{100 * v[0] / tot:5.1f}% smp {100 * v[1] / toti:5.1f}% inst  {k[0]}:{k[1]:<4d} {v[2]}")

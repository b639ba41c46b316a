"""Aggregate an `ncu --page source --csv` dump by code region.  Usage: ncu_regions.py dump.csv nworld regions.txt
regions.txt lines: file first_line last_line name"""
import csv
import sys
from collections import defaultdict

agg = defaultdict(lambda: [0, 0])
fname = ""; hdr = None
for r in csv.reader(open(sys.argv[1])):
  if not r:
    continue
  if r[0] == "File Path":
    fname = r[1].split("/")[-1]; continue
  if r[0] == "Function Name":
    continue
  if r[0] == "Line No":
    hdr = r; si = hdr.index("# Samples"); ii = hdr.index("Instructions Executed"); continue
  if hdr is None or r[2] != "-":
    continue
  try:
    s = int(r[si]); n = int(r[ii])
  except ValueError:
    continue
  agg[(fname, int(r[0]))][0] += s; agg[(fname, int(r[0]))][1] += n
W = int(sys.argv[2])
regions = [l.split() for l in open(sys.argv[3]) if l.strip() and not l.startswith("#")]
tot = sum(v[0] for v in agg.values()) or 1; toti = sum(v[1] for v in agg.values()) or 1
out = defaultdict(lambda: [0, 0])
for (f, l), v in agg.items():
  name = "other:" + f
  for rf, a, b, n in regions:
    if f == rf and int(a) <= l <= int(b):
      name = n
  out[name][0] += v[0]; out[name][1] += v[1]
print(f"total samples {tot}, warp-instructions {toti} ({toti / W:.0f} per world)")
for k, v in sorted(out.items(), key=lambda kv: -kv[1][0]):
  print(f"{k:28s} {100 * v[0] / tot:5.1f}% smp {100 * v[1] / toti:5.1f}% inst  {v[1] / W:8.0f} inst/world")

"""Attribute the samples of an `ncu --page source --csv --print-source cuda,sass` dump to the nearest preceding line of ONE source
file in SASS address order (inlined helpers are charged to their call site).  Usage: ncu_callsite.py dump.csv file.cu [top] [bucket]"""
import csv, sys
from collections import defaultdict
main = sys.argv[2]; top = int(sys.argv[3]) if len(sys.argv) > 3 else 40; bucket = int(sys.argv[4]) if len(sys.argv) > 4 else 1
ins = []   # (addr, file, line, samples, inst)
fname = ""; line = 0; hdr = None
for r in csv.reader(open(sys.argv[1])):
  if not r: continue
  if r[0] == "File Path": fname = r[1].split("/")[-1]; continue
  if r[0] in ("Function Name",): continue
  if r[0] == "Line No":
    hdr = r; si = hdr.index("# Samples"); ii = hdr.index("Instructions Executed")
    st = [(k, h[6:]) for k, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]; continue
  if hdr is None: continue
  if r[2] == "-": line = int(r[0]); src = r[1]; continue
  if not r[2].startswith("0x"): continue
  try: ins.append((int(r[2], 16), fname, line, int(r[si]), int(r[ii]), [int(r[k] or 0) for k, _ in st]))
  except ValueError: pass
ins.sort()
agg = defaultdict(lambda: [0, 0, None]); cur = 0
for a, f, l, s, n, sv in ins:
  if f == main: cur = l
  e = agg[cur // bucket * bucket]
  e[0] += s; e[1] += n; e[2] = sv if e[2] is None else [x + y for x, y in zip(e[2], sv)]
tot = sum(v[0] for v in agg.values()) or 1; toti = sum(v[1] for v in agg.values()) or 1
src = open([p for p in ("asimov-mjlab_b200/csrc/" + main,) ][0]).read().splitlines()
print(f"total samples {tot} inst {toti}")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
  print(f"{100*v[0]/tot:5.1f}% smp {100*v[1]/toti:5.1f}% inst  {main}:{k:<4d} {src[k-1].strip()[:70] if 0 < k <= len(src) else ''}  | " +
        " ".join(f"{nm}:{c}" for c, nm in sorted(((c, nm) for c, (_, nm) in zip(v[2], st)), reverse=True)[:3] if c))

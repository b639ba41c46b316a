"""Static SASS instruction count per source region of one kernel.
Usage: sass_regions.py file.sass kernel_substring regions.txt   (file.sass from `nvdisasm -g -c x.cubin`)
regions.txt lines: file first_line last_line name"""
import collections
import re
import sys

regions = [l.split() for l in open(sys.argv[3]) if l.strip() and not l.startswith("#")]
cur = None; fn = None
cnt = collections.Counter()
for l in open(sys.argv[1]):
  m = re.match(r"\s*\.text\.(\S+):", l)
  if m:
    fn = m.group(1); continue
  m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
  if m:
    cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
  if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l) and fn and sys.argv[2] in fn and cur:
    name = "other:" + cur[0]
    for rf, a, b, n in regions:
      if cur[0] == rf and int(a) <= cur[1] <= int(b):
        name = n
    cnt[name] += 1
tot = sum(cnt.values())
print(f"{sys.argv[2]}: {tot} SASS instructions ({tot * 16 / 1024:.0f} KB)")
for k, v in cnt.most_common():
  print(f"{k:28s} {v:6d}  {100 * v / tot:5.1f}%")

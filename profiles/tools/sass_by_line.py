"""joins the SASS page of an ncu report (`ncu -i rep --page source --csv`) with `nvdisasm -g -c`
of the same cubin: stall samples and executed instructions per source line / per phase marker.
    python profiles/tools/sass_by_line.py sass_page.csv kernel.sass [source.cuh]"""
import csv
import re
import sys
from collections import defaultdict

rows = list(csv.reader(open(sys.argv[1])))
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hdr]
si, ii = h.index("# Samples"), h.index("Instructions Executed")
prof = []
for r in rows[hdr + 1:]:  # (the first kernel of the report)
    if not r or r[0] in ("Kernel Name", "Address"):
        break
    if len(r) > ii:
        prof.append((float(r[si] or 0), float(r[ii] or 0)))
line, lines = None, []
for l in open(sys.argv[2]):
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        line = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
        lines.append(line)
print(len(prof), "profiled instructions,", len(lines), "disassembled")
n = min(len(prof), len(lines))
by = defaultdict(lambda: [0.0, 0.0])
for (s, i), ln in zip(prof[:n], lines[:n]):
    by[ln][0] += s
    by[ln][1] += i
ts, ti = sum(v[0] for v in by.values()), sum(v[1] for v in by.values())
src = {}
if len(sys.argv) > 3:
    src = {k + 1: t.rstrip() for k, t in enumerate(open(sys.argv[3]))}
for ln, (s, i) in sorted(by.items(), key=lambda kv: -kv[1][0])[:40]:
    text = src.get(ln[1], "") if ln and ln[0] == sys.argv[3].split("/")[-1] else ""
    print(f"{str(ln):34s} {100 * s / ts:5.1f}% samples {100 * i / ti:5.1f}% instr  {text.strip()[:90]}")

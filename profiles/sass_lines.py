"""Attributes an ncu `--page source --csv` (SASS) dump to CUDA source lines through nvdisasm -g line info.

usage: python profiles/sass_lines.py <ncu source csv> <nvdisasm -g output> <kernel substring> [top N]

  ncu -i X.ncu-rep --page source --csv > src.csv
  cuobjdump -xelf all cnf2freq_b200/lib/libcnf2_engine.so && nvdisasm -g *.cubin > all.sass
"""
import csv
import re
import sys
from collections import defaultdict

src_csv, sass, kern = sys.argv[1:4]
topn = int(sys.argv[4]) if len(sys.argv) > 4 else 40
# nvdisasm: find the kernel's .text section, walk instructions in order with the current (file, line)
lines = open(sass).read().split("\n")
start = next(i for i, l in enumerate(lines) if l.startswith("\t.section\t.text.") and kern in l)
cur = ("?", 0)
order = []
for l in lines[start + 1:]:
    if l.startswith("\t.section"):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", l):
        order.append(cur)
rows = list(csv.reader(open(src_csv)))
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
H = rows[hdr]
data = [dict(zip(H, r)) for r in rows[hdr + 1:] if len(r) == len(H)]
assert len(data) == len(order), (len(data), len(order))
agg = defaultdict(lambda: [0.0, 0.0])
tot = [0.0, 0.0]
for d, where in zip(data, order):
    inst = float(d["Instructions Executed"] or 0)
    smp = float(d["# Samples"] or 0)
    a = agg[where]
    a[0] += inst
    a[1] += smp
    tot[0] += inst
    tot[1] += smp
print(f"total warp instructions {tot[0]:.3e}, samples {tot[1]:.0f}")
srcs = {}
for (f, ln), (inst, smp) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:topn]:
    if f not in srcs:
        try:
            srcs[f] = open("cnf2freq_b200/csrc/" + f).read().split("\n")
        except OSError:
            srcs[f] = []
    text = srcs[f][ln - 1].strip()[:100] if 0 < ln <= len(srcs[f]) else ""
    print(f"{f[:22]:22s} {ln:4d}  inst {100 * inst / tot[0]:5.1f}%  samples {100 * smp / tot[1]:5.1f}%  {text}")

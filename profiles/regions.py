"""Aggregates an ncu `--page source --csv` dump by source FUNCTION (from nvdisasm -g line info mapped onto
the `__device__`/`__global__` function that contains each line) and prints stall-reason totals.

usage: python profiles/regions.py <report.ncu-rep> <kernel substring>
"""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict

rep, kern = sys.argv[1:3]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "cnf2freq_b200/lib/libcnf2_engine.so")], cwd=tmp, stdout=subprocess.DEVNULL)
cubins = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")]
sass = subprocess.run(["nvdisasm", "-g"] + cubins, capture_output=True, text=True).stdout.split("\n")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
start = next(i for i, l in enumerate(sass) if l.startswith("\t.section\t.text.") and kern in l)
cur, order = ("?", 0), []
for l in sass[start + 1:]:
    if l.startswith("\t.section"):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", l):
        order.append(cur)
rows = list(csv.reader(io.StringIO(src)))
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
H = rows[hdr]
data = [dict(zip(H, r)) for r in rows[hdr + 1:] if len(r) == len(H)]
assert len(data) == len(order), (len(data), len(order))
funcs = {}


def func_of(f, ln):
    if f not in funcs:
        tab = []
        try:
            for i, l in enumerate(open(os.path.join(ROOT, "cnf2freq_b200/csrc", f)).read().split("\n"), 1):
                m = re.match(r"^(?:template.*>\s*)?(?:__device__|__global__|__host__).*?\b(\w+)\s*\(", l)
                if m:
                    tab.append((i, m.group(1)))
        except OSError:
            pass
        funcs[f] = tab
    name = f
    for i, n in funcs[f]:
        if i <= ln:
            name = n
    return name


agg = defaultdict(lambda: [0.0, 0.0])
tot = [0.0, 0.0]
for d, (f, ln) in zip(data, order):
    i, s = float(d["Instructions Executed"] or 0), float(d["# Samples"] or 0)
    r = func_of(f, ln)
    agg[r][0] += i
    agg[r][1] += s
    tot[0] += i
    tot[1] += s
print(f"total warp instructions {tot[0]:.4e}, samples {tot[1]:.0f}")
for r, (i, s) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{r:28s} inst {100 * i / tot[0]:5.1f}%  samples {100 * s / tot[1]:5.1f}%")
raw = list(csv.reader(io.StringIO(subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout)))
vals = dict(zip(raw[0], raw[2]))
st = {k.split("issue_stalled_")[1]: float(v) for k, v in vals.items() if "pcsamp_warps_issue_stalled" in k and "not_issued" not in k}
ts = sum(st.values())
print("stalls: " + ", ".join(f"{k} {100 * v / ts:.0f}%" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:8]))
for k in ("gpu__time_duration.sum", "smsp__inst_executed.sum", "sm__inst_executed.avg.per_cycle_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
          "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__occupancy_limit_shared_mem"):
    if k in vals:
        print(k, vals[k])

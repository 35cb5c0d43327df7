#!/usr/bin/env python
"""Per-kernel registers / spills / stack from a build log (-Xptxas -v).  usage: ptxas_report.py [log] [filter]"""
import re, subprocess, sys
log = sys.argv[1] if len(sys.argv) > 1 else "sparc_b200/_lib/libsparc_b200.so.build.log"
flt = sys.argv[2] if len(sys.argv) > 2 else ""
txt = open(log).read().splitlines()
cur = None
rows = {}
for ln in txt:
    m = re.search(r"Compiling entry function '(\S+)'", ln)
    if m:
        cur = m.group(1); rows[cur] = {}
        continue
    m = re.search(r"Function properties for (\S+)", ln)
    if m:
        cur = m.group(1); rows.setdefault(cur, {})
        continue
    if cur is None:
        continue
    m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", ln)
    if m:
        rows[cur].update(stack=int(m.group(1)), sst=int(m.group(2)), sld=int(m.group(3)))
    m = re.search(r"Used (\d+) registers", ln)
    if m:
        rows[cur]["regs"] = int(m.group(1))
names = list(rows)
dem = subprocess.run(["c++filt"] + names, capture_output=True, text=True).stdout.splitlines()
for n, d in zip(names, dem):
    if flt and flt not in d:
        continue
    r = rows[n]
    d = re.sub(r"\(SparcParams.*", "", d)
    print(f"{r.get('regs','-'):>4} regs  stack {r.get('stack','-'):>4}  spill st/ld {r.get('sst','-'):>4}/{r.get('sld','-'):>4}  {d}")

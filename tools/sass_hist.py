"""Summarise an `ncu --page source --csv --print-source sass` export: executed warp-instructions
per env-µs by opcode, and the hottest source ranges.  usage: sass_hist.py src.csv n_env_us"""
import csv, sys
from collections import Counter
rows = list(csv.reader(open(sys.argv[1])))
n_env_us = float(sys.argv[2]) if len(sys.argv) > 2 else 4.096e6
hdr = rows[1]
iS, iN, iSamp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
ins = []
for r in rows[2:]:
    if r and r[0] == "Kernel Name":
        break
    if len(r) < len(hdr) or r[0] == "Address":
        continue
    ins.append((r[iS].strip(), int(r[iN]), int(r[iSamp])))
tot = sum(n for _, n, _ in ins)
print(f"total warp-inst {tot}  ({tot / n_env_us:.1f} per env-us), {len(ins)} SASS lines")
c, cs = Counter(), Counter()
for s, n, sm in ins:
    parts = s.split()
    op = parts[1] if parts[0].startswith('@') else parts[0]
    op = op.split('.')[0]
    c[op] += n
    cs[op] += sm
for op, n in c.most_common(28):
    print(f"{op:10s} {n / n_env_us:8.2f} per env-us   stall samples {cs[op]}")

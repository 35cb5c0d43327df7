"""Aggregate an ncu SASS source-page export by CUDA source line.

ncu's CSV export of the source page carries per-SASS-instruction metrics only; `nvdisasm -g -c` of the
same cubin carries the line table.  Both list the kernel's instructions in address order, so they are
joined by position.
usage:
  cuobjdump -xelf all sparc_b200/_lib/libsparc_b200.so.obj/step_inst_w1.o        # -> *.cubin
  nvdisasm -g -c step_inst_w1.sm_100a.cubin > dis.txt
  ncu -i rep.ncu-rep --page source --csv --kernel-name regex:step_kernel > src.csv
  python tools/line_profile.py dis.txt src.csv <mangled-kernel-substring> <n_env_us>
"""
import collections
import csv
import re
import sys

dis, src, kern, n_env_us = sys.argv[1], sys.argv[2], sys.argv[3], float(sys.argv[4])
seq, cur, on = [], None, False
for l in open(dis):
    if ".text." in l and l.lstrip().startswith("//-----"):
        on = kern in l
        continue
    if not on:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s*/\*[0-9a-f]{4,6}\*/\s+\S", l):
        seq.append(cur)
rows = list(csv.reader(open(src)))
hdr, data = rows[1], rows[2:]
isamp, iex = hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
assert len(seq) == len(data), (len(seq), len(data))
tot = sum(int(r[isamp] or 0) for r in data)
agg = collections.defaultdict(lambda: [0, 0])
for where, r in zip(seq, data):
    agg[where][0] += int(r[iex] or 0)
    agg[where][1] += int(r[isamp] or 0)
print(f"{len(data)} SASS instructions, {sum(v[0] for v in agg.values()) / n_env_us:.1f} warp-inst per env-us")
for (f, ln), (ex, s) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:40]:
    print(f"{f}:{ln:4d}  {ex / n_env_us:6.2f} inst/env-us  {100 * s / tot:5.1f} % of stall samples")

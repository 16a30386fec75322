"""Summarise an `ncu --page source --csv` export: instruction mix, stall mix, hottest SASS lines.

    ncu -i X.ncu-rep --page source --csv --kernel-name regex:NAME --launch-count 1 > src.csv
    python profiles/sass_summary.py src.csv [top]
"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 20
h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[h]
data = [r for r in rows[h + 1 :] if len(r) == len(hdr) and r[0] != "Address"]
num = lambda r, name: int(float(r[hdr.index(name)] or 0))
iSrc = hdr.index("Source")
S, I = "Warp Stall Sampling (All Samples)", "Instructions Executed"
tot, toti = sum(num(r, S) for r in data), sum(num(r, I) for r in data)
print(f"kernel: {rows[0][1][:80]}\nsamples {tot}  warp-instructions {toti}  sass lines {len(data)}")
op_s, op_i = collections.Counter(), collections.Counter()
for r in data:
    s = r[iSrc].split()
    o = (s[1] if s[0].startswith("@") else s[0]).split(".")[0]
    op_s[o] += num(r, S)
    op_i[o] += num(r, I)
for o, c in op_i.most_common(top):
    print(f"  {o:10s} instr {c / max(toti, 1) * 100:5.1f}%   stall samples {op_s[o] / max(tot, 1) * 100:5.1f}%")
stalls = [x for x in hdr if x.startswith("stall_") and "Not" not in x]
agg = sorted(((sum(num(r, s) for r in data), s) for s in stalls), reverse=True)
print("stall reasons:", ", ".join(f"{s[6:]} {c / max(tot, 1) * 100:.1f}%" for c, s in agg[:8]))
print("hottest lines:")
for r in sorted(data, key=lambda r: -num(r, S))[:top]:
    print(f"  {r[0][-5:]}  {num(r, S) / max(tot, 1) * 100:5.1f}%  x{num(r, I):>10d}  {r[iSrc][:100]}")

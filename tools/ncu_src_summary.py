"""Summarise an `ncu --page source --csv` dump: instruction mix by opcode and hottest lines.
usage: python tools/ncu_src_summary.py file.csv [top_n]"""
import collections, csv, re, sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hi = next(i for i, r in enumerate(rows) if len(r) > 5 and r[1] == "Source")
hdr = rows[hi]
iA, iS, iI = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
ops, samp, lines = collections.Counter(), collections.Counter(), []
tot_i = tot_s = 0
for r in rows[hi + 1:]:
    if len(r) < len(hdr) or not r[iI].isdigit():
        continue
    ins, s = int(r[iI]), int(r[iS] or 0)
    m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[iA])
    op = m.group(2).split(".")[0] if m else "?"
    ops[op] += ins
    samp[op] += s
    tot_i += ins
    tot_s += s
    lines.append((s, ins, r[iA].strip()[:90]))
print("total warp-instructions", tot_i, "stall samples", tot_s)
for op, c in ops.most_common(top):
    print(f"{op:12s} {c:12d} {100*c/tot_i:5.1f}%   samples {100*samp[op]/max(tot_s,1):5.1f}%")
print("--- hottest instructions by stall samples")
for s, ins, src in sorted(lines, reverse=True)[:top]:
    print(f"{s:7d} {ins:10d}  {src}")

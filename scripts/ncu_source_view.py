"""Read `ncu -i X.ncu-rep --page source --csv -k regex:K` (stdin or file): stall-reason totals, opcode histogram, hottest lines."""
import collections
import csv
import re
import sys

path = sys.argv[1]
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 30
rows = list(csv.reader(open(path)))
hdr, data = rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
S, I = ix["Warp Stall Sampling (All Samples)"], ix["Instructions Executed"]
cols = [(h, i) for h, i in ix.items() if h.startswith("stall_") and "Not Issued" not in h]
tot = collections.Counter()
for r in data:
    for h, i in cols:
        if r[i]:
            tot[h] += int(r[i])
n = sum(tot.values())
print("total samples", n, "warp-instructions", sum(int(r[I] or 0) for r in data))
for h, c in tot.most_common(10):
    print(f"  {h:26s}{c:8d} {100 * c / n:5.1f}%")
op_i, op_s = collections.Counter(), collections.Counter()
for r in data:
    m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[ix["Source"]])
    op = m.group(2).split(".")[0] if m else "?"
    op_i[op] += int(r[I] or 0)
    op_s[op] += int(r[S] or 0)
ti, ts = sum(op_i.values()), sum(op_s.values())
print("--- opcode: instr% sample%")
for op, c in op_i.most_common(14):
    print(f"  {op:12s}{100 * c / ti:6.2f}% {100 * op_s[op] / ts:6.2f}%")
print("--- hottest lines")
for r in sorted(data, key=lambda r: -int(r[S] or 0))[:topn]:
    reasons = sorted(((h.replace("stall_", ""), int(r[i])) for h, i in cols if r[i] not in ("", "0")), key=lambda x: -x[1])
    print(r[S].rjust(6), r[I].rjust(9), r[ix["Source"]].strip()[:72].ljust(72), reasons[:2])

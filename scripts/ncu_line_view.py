"""Per-source-line stall samples of one kernel: joins `ncu --page source --csv` (SASS rows with addresses) with the line
table of the cubin that ran (`nvdisasm -g`, extracted from the in-tree .so with `cuobjdump -xelf`).

    python scripts/ncu_line_view.py <sass.csv> <kernel-substring> [top-n] [so-path]

The .so must be the build the report was taken from.  Inlined code is attributed to the innermost line nvdisasm prints
before the instruction (the `//## File ..., line N inlined at ...` chain's first entry)."""
import collections
import csv
import glob
import os
import re
import subprocess
import sys
import tempfile

csv_path, kernel = sys.argv[1], sys.argv[2]
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 25
so = sys.argv[4] if len(sys.argv) > 4 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                         "distriflow_b200", "ops", "libdistriflow_b200.so")
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=tmp, capture_output=True)
line_of = {}
for cubin in glob.glob(os.path.join(tmp, "*.cubin")):
    txt = subprocess.run(["nvdisasm", "-g", cubin], capture_output=True, text=True).stdout
    inside, cur = False, None
    for ln in txt.splitlines():
        if ln.startswith("\t.section\t.text."):
            inside = kernel in ln
            cur = None
            continue
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (m.group(1), int(m.group(2)))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+\S", ln)
        if m and ".byte" not in ln:
            line_of[int(m.group(1), 16)] = cur
    if line_of:
        break
if not line_of:
    sys.exit(f"kernel {kernel!r} not found in {so}")
rows = list(csv.reader(open(csv_path)))
hdr, data = rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
S, I = ix["Warp Stall Sampling (All Samples)"], ix["Instructions Executed"]
cols = [(h, i) for h, i in ix.items() if h.startswith("stall_") and "Not Issued" not in h]
base = min(int(r[ix["Address"]], 16) for r in data)
samples, instrs, why = collections.Counter(), collections.Counter(), collections.defaultdict(collections.Counter)
for r in data:
    off = int(r[ix["Address"]], 16) - base
    ln = line_of.get(off)
    samples[ln] += int(r[S] or 0)
    instrs[ln] += int(r[I] or 0)
    for h, i in cols:
        if r[i] not in ("", "0"):
            why[ln][h.replace("stall_", "")] += int(r[i])
cache = {}


def text_of(key):
    if key is None:
        return "?"
    f, n = key
    if f not in cache:
        cache[f] = open(f).read().splitlines() if os.path.exists(f) else []
    return cache[f][n - 1].strip()[:84] if n <= len(cache[f]) else "?"


tot_s, tot_i = sum(samples.values()), sum(instrs.values())
print(f"{kernel}: {tot_s} samples, {tot_i} warp-instructions")
print("samples  share   instr%  file:line  source")
for key, c in samples.most_common(topn):
    where = f"{os.path.basename(key[0])}:{key[1]}" if key else "?"
    top = ", ".join(f"{k} {v}" for k, v in why[key].most_common(3))
    print(f"{c:7d} {100 * c / tot_s:5.1f}%  {100 * instrs[key] / max(tot_i, 1):5.1f}%  {where}  {text_of(key)}\n{'':24s}[{top}]")

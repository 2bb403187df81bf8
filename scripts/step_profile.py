"""Per-kernel share of one training step, from an `ncu --metrics gpu__time_duration.sum --csv` launch list.

Usage (on the GPU box, after the same command has exited 0 without ncu):

    python bench.py --steps 2 --warmup 3 --no-graph --no-e2e
    DFB_PROFILE_RANGE=1 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
        --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-graph --no-e2e
    python scripts/step_profile.py gpurun_out/launches.csv --steps 2 > gpurun_out/step_profile.md

`DFB_PROFILE_RANGE=1` makes bench.py bracket its timed steps with cudaProfilerStart/Stop, so the capture holds exactly
`--steps` identical steps.  Kernel times under ncu are serialised and cold-cache: read the shares, not the sum.
"""
from __future__ import annotations

import argparse
import collections
import csv
import re
import sys


def short(name: str) -> str:
    name = re.sub(r"^void\s+", "", name)
    name = re.sub(r"\(anonymous namespace\)::", "", name)
    m = re.match(r"([A-Za-z0-9_:]+)", name)
    base = m.group(1) if m else name
    if "vectorized_elementwise_kernel" in name or "elementwise_kernel" in base:
        f = re.search(r"native::([A-Za-z0-9_]+(?:Functor|kernel_cuda|_kernel)[A-Za-z0-9_]*)", name)
        return "aten elementwise: " + (f.group(1) if f else "?")
    return base.split("::")[-1] if base.startswith(("at::", "native::")) else base


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("--steps", type=int, default=1, help="identical steps contained in the capture")
    ap.add_argument("--skip", type=int, default=0, help="launches to drop from the head (model construction)")
    ap.add_argument("--top", type=int, default=24)
    a = ap.parse_args()
    rows = []
    with open(a.csv, newline="") as f:
        lines = [ln for ln in f if ln.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") == "gpu__time_duration.sum":
            rows.append((r["Kernel Name"], float(r["Metric Value"].replace(",", ""))))
    rows = rows[a.skip:]
    agg = collections.OrderedDict()
    for name, ns in rows:
        k = short(name)
        c = agg.setdefault(k, [0, 0.0])
        c[0] += 1
        c[1] += ns
    total = sum(v[1] for v in agg.values())
    print(f"launches {len(rows)} ({len(rows) / a.steps:.0f} per step), kernel time {total / a.steps / 1e6:.2f} ms per step "
          f"(serialised under ncu)\n")
    print("| kernel | launches/step | ms/step | share |")
    print("|---|---|---|---|")
    for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:a.top]:
        print(f"| `{k}` | {n / a.steps:.1f} | {ns / a.steps / 1e6:.3f} | {100 * ns / total:.1f} % |")
    return 0


if __name__ == "__main__":
    sys.exit(main())

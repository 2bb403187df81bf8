"""SASS mnemonic counts per kernel of the built library (runs on the build machine: cuobjdump needs no GPU).

    python scripts/sass_summary.py > profiles/sass_summary_r1_final.txt

UTCHMMA / UTCQMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG / UTMASTG / UTMAREDG = TMA load / store / reduce,
UTCBAR = tcgen05.commit, SYNCS = mbarrier, LDGMC = multimem.ld_reduce, `.SYS` = system-scope (peer-memory) accesses.
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distriflow_b200", "ops", "libdistriflow_b200.so")
WANT = ("UTCHMMA", "UTCQMMA", "UTCBAR", "UTCATOMSWS", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UTMAPF", "SYNCS", "LDGMC",
        "REDG", "ATOMG", "UCGABAR")


def main() -> int:
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    kernels, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = kernels.setdefault(m.group(1), collections.Counter())
            continue
        if cur is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
        if not m:
            continue
        op = m.group(1)
        base = op.split(".")[0]
        if base in WANT:
            cur[base] += 1
        if ".SYS" in op and base in ("LDG", "LD"):
            cur["sys_loads"] += 1
        if ".SYS" in op and base in ("STG", "ST", "REDG", "RED", "ATOMG"):
            cur["sys_stores"] += 1
    names = subprocess.run(["c++filt"], input="\n".join(kernels), capture_output=True, text=True).stdout.splitlines()
    print(__doc__.strip().split("\n\n")[-1])
    print(f"# {len(kernels)} kernels in {os.path.relpath(LIB, ROOT)}\n")
    for name, (_, c) in zip(names, kernels.items()):
        print(name[:150])
        print("    " + (", ".join(f"{k}={v}" for k, v in c.items()) or "(CUDA-core kernel: no tensor / TMA / mbarrier instructions)"))
    return 0


if __name__ == "__main__":
    sys.exit(main())

"""Build the native library in-tree: every `ops/csrc/*.cu` and `parallel/csrc/*.cu` -> `ops/libdistriflow_b200.so`.

`python -m distriflow_b200.build [--force] [-j N]`.  nvcc cross-compiles for sm_100a without a GPU; the
result travels to the GPU box inside the source tree.  Objects are cached under `build/` keyed by a hash
of the source + headers + flags, so an unchanged file is not recompiled.
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import glob
import hashlib
import os
import shutil
import subprocess
import sys
from typing import List

ROOT = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(ROOT)
SRC_DIRS = [os.path.join(ROOT, "ops", "csrc"), os.path.join(ROOT, "parallel", "csrc")]
OUT = os.path.join(ROOT, "ops", "libdistriflow_b200.so")
BUILD_DIR = os.path.join(REPO, "build", "obj")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--use_fast_math",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "--expt-relaxed-constexpr"]


def nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found")
    return exe


def sources() -> List[str]:
    out: List[str] = []
    for d in SRC_DIRS:
        out += sorted(glob.glob(os.path.join(d, "*.cu")))
    return out


def _headers_hash() -> str:
    h = hashlib.sha256()
    for d in SRC_DIRS + [os.path.join(REPO, "include")]:
        for f in sorted(glob.glob(os.path.join(d, "*.cuh")) + glob.glob(os.path.join(d, "*.h"))):
            h.update(open(f, "rb").read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def _compile(src: str, hh: str, verbose: bool) -> str:
    key = hashlib.sha256(open(src, "rb").read() + hh.encode()).hexdigest()[:16]
    obj = os.path.join(BUILD_DIR, f"{os.path.basename(src)}.{key}.o")
    if os.path.exists(obj):
        return obj
    cmd = [nvcc(), *NVCC_FLAGS, "-I", os.path.join(REPO, "include"), *sum((["-I", d] for d in SRC_DIRS), []),
           "-c", src, "-o", obj]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    if verbose:
        sys.stderr.write(r.stderr)
    return obj


def build(force: bool = False, jobs: int = 8, verbose: bool = False) -> str:
    os.makedirs(BUILD_DIR, exist_ok=True)
    srcs = sources()
    hh = _headers_hash()
    if force:
        for f in glob.glob(os.path.join(BUILD_DIR, "*.o")):
            os.remove(f)
    with cf.ThreadPoolExecutor(max_workers=jobs) as ex:
        objs = list(ex.map(lambda s: _compile(s, hh, verbose), srcs))
    stamp = hashlib.sha256(" ".join(sorted(objs)).encode()).hexdigest()
    stamp_file = OUT + ".stamp"
    if not force and os.path.exists(OUT) and os.path.exists(stamp_file) and open(stamp_file).read() == stamp:
        return OUT
    cmd = [nvcc(), "-shared", "-o", OUT + ".tmp", *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-lcudart_static",
           "-lpthread", "-ldl", "-lrt"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    os.replace(OUT + ".tmp", OUT)
    with open(stamp_file, "w") as f:
        f.write(stamp)
    keep = set(objs)                                    # drop objects of older source revisions from the cache
    for f in glob.glob(os.path.join(BUILD_DIR, "*.o")):
        if f not in keep:
            os.remove(f)
    return OUT


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("-j", type=int, default=8)
    ap.add_argument("-v", action="store_true")
    a = ap.parse_args(argv)
    print(build(a.force, a.j, a.v))
    return 0


if __name__ == "__main__":
    sys.exit(main())

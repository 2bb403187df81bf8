"""`include/distriflow_b200.h` declares every `extern "C"` entry point of the native library, and nothing else.

No reference counterpart (the reference has no native code); this keeps the published C ABI honest."""
import glob
import os
import re
import subprocess
import shutil

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _defined():
    names = set()
    for f in glob.glob(os.path.join(ROOT, "distriflow_b200", "*", "csrc", "*.cu")):
        for m in re.finditer(r"DFB_API\s+[\w\s\*]+?\b(dfb_\w+)\s*\(", open(f).read()):
            names.add(m.group(1))
    return names


def test_header_declares_every_entry_point():
    hdr = open(os.path.join(ROOT, "include", "distriflow_b200.h")).read()
    declared = set(re.findall(r"\b(dfb_\w+)\s*\(", hdr))
    defined = _defined()
    assert len(defined) > 40
    assert defined - declared == set(), f"missing from the header: {sorted(defined - declared)}"
    assert declared - defined == set(), f"declared but not defined: {sorted(declared - defined)}"


@pytest.mark.skipif(shutil.which("gcc") is None, reason="no C compiler")
def test_header_is_plain_c(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "distriflow_b200.h"\nint main(void) { return 0; }\n')
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-c", str(src), "-I", os.path.join(ROOT, "include"), "-o",
                    str(tmp_path / "t.o")], check=True)

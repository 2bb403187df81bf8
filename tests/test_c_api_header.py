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


def _header_params():
    """{name: number of parameters} from the declarations in the public header."""
    hdr = open(os.path.join(ROOT, "include", "distriflow_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", " ", hdr, flags=re.S)
    out = {}
    for m in re.finditer(r"\b(dfb_\w+)\s*\(([^;{]*?)\)\s*;", hdr, flags=re.S):
        params = m.group(2).strip()
        out[m.group(1)] = 0 if params in ("", "void") else len([p for p in params.split(",") if p.strip()])
    return out


def test_ctypes_signatures_match_the_header():
    """Every Python wrapper's `argtypes` has as many entries as the C declaration has parameters: a signature changed on
    one side only would otherwise corrupt arguments silently (the library loads without a GPU, so this runs on CPU)."""
    import ctypes
    from distriflow_b200.ops import _lib
    if not _lib.available():
        pytest.skip("native library not built")
    from distriflow_b200.ops import attention, conv, dense, norm, softmax_ce
    from distriflow_b200.parallel import engine, symm
    L = _lib.lib()
    for init in (attention._sigs, conv._sigs, dense._init_sigs, norm._sigs, softmax_ce._sigs, engine._comm_sigs, symm._sigs):
        init()
    decl = _header_params()
    checked = 0
    for name, nparams in decl.items():
        fn = getattr(L, name)                          # AttributeError here = declared but not exported
        if fn.argtypes is None:
            continue
        assert len(fn.argtypes) == nparams, f"{name}: ctypes has {len(fn.argtypes)} arguments, the header {nparams}"
        checked += 1
    assert checked >= 35, checked
    assert L.dfb_last_error.restype is ctypes.c_char_p


def _kind_of_c(param: str) -> str:
    p = param.strip()
    if "*" in p:
        return "ptr"
    words = p.split()
    if "float" in words:
        return "f32"
    if "size_t" in words or "long" in words:
        return "i64"                 # size_t / long / unsigned long long: one 64-bit integer register on LP64
    return "i32"                     # int / unsigned


def _kind_of_ctype(t) -> str:
    import ctypes
    if t in (ctypes.c_void_p, ctypes.c_char_p) or isinstance(t, type) and issubclass(t, ctypes._Pointer):
        return "ptr"
    if t is ctypes.c_float:
        return "f32"
    if t in (ctypes.c_size_t, ctypes.c_ulonglong, ctypes.c_long, ctypes.c_int64, ctypes.c_ulong):
        return "i64"
    return "i32" if t in (ctypes.c_int, ctypes.c_uint) else f"?{t}"


def test_ctypes_argument_kinds_match_the_header():
    """Position by position: pointer / 32-bit integer / 64-bit integer / float agree between the header and the
    ctypes tables (a float passed where the C side reads an int is the bug this guards against)."""
    from distriflow_b200.ops import _lib
    if not _lib.available():
        pytest.skip("native library not built")
    from distriflow_b200.ops import attention, conv, dense, norm, softmax_ce
    from distriflow_b200.parallel import engine, symm
    L = _lib.lib()
    for init in (attention._sigs, conv._sigs, dense._init_sigs, norm._sigs, softmax_ce._sigs, engine._comm_sigs, symm._sigs):
        init()
    hdr = re.sub(r"/\*.*?\*/", " ", open(os.path.join(ROOT, "include", "distriflow_b200.h")).read(), flags=re.S)
    bad = []
    for m in re.finditer(r"\b(dfb_\w+)\s*\(([^;{]*?)\)\s*;", hdr, flags=re.S):
        fn = getattr(L, m.group(1))
        if fn.argtypes is None or m.group(2).strip() in ("", "void"):
            continue
        for i, (cp, ct) in enumerate(zip(m.group(2).split(","), fn.argtypes)):
            kc, kt = _kind_of_c(cp), _kind_of_ctype(ct)
            if kc != kt:
                bad.append(f"{m.group(1)} arg {i} ({cp.strip()}): header {kc}, ctypes {kt}")
    assert not bad, "\n".join(bad)


def test_header_matches_the_definitions_parameter_by_parameter():
    """The declarations in the header and the `DFB_API` definitions in the .cu sources have the same parameter kinds in the
    same order (cudaStream_t on the definition side is the header's `void* stream`)."""
    hdr = re.sub(r"/\*.*?\*/", " ", open(os.path.join(ROOT, "include", "distriflow_b200.h")).read(), flags=re.S)
    decl = {m.group(1): m.group(2) for m in re.finditer(r"\b(dfb_\w+)\s*\(([^;{]*?)\)\s*;", hdr, flags=re.S)}
    bad, seen = [], 0
    for f in glob.glob(os.path.join(ROOT, "distriflow_b200", "*", "csrc", "*.cu")):
        src = re.sub(r"//[^\n]*", " ", open(f).read())
        src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)
        for m in re.finditer(r"DFB_API\s+[\w\s\*]+?\b(dfb_\w+)\s*\(([^{;]*?)\)\s*\{", src, flags=re.S):
            name, params = m.group(1), m.group(2)
            seen += 1

            def kinds(plist):
                plist = plist.strip()
                if plist in ("", "void"):
                    return []
                return ["ptr" if "cudaStream_t" in p else _kind_of_c(p) for p in plist.split(",")]

            if kinds(params) != kinds(decl[name]):
                bad.append(f"{name}: definition {kinds(params)} vs header {kinds(decl[name])}")
    assert seen > 40 and not bad, "\n".join(bad)

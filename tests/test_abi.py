"""The C-ABI library loads and exports every symbol include/stitch.h declares (no GPU needed)."""
import ctypes as C
import os
import re

from imagestitching_b200 import _lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "stitch.h"), encoding="utf-8").read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = set(re.findall(r"\b(stitch_[a-z0-9_]+)\s*\(", text))
    return names


def test_every_declared_symbol_is_exported_and_bound():
    lib = L.load()
    declared = _declared()
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in stitch.h but not exported"
    assert declared == set(L.SIGNATURES), declared ^ set(L.SIGNATURES)


def test_abi_version_and_strerror():
    lib = L.load()
    assert lib.stitch_abi_version() == 8
    assert lib.stitch_strerror(L.E_BAD_DIMS).decode() == "图片尺寸不可用"
    assert lib.stitch_strerror(L.E_BAD_OUT).decode() == "输出尺寸计算失败"
    assert lib.stitch_strerror(L.E_NO_SURFACE).decode() == "无法获取绘图上下文"


def test_struct_sizes_match_header():
    # sizes computed by hand from include/stitch.h (LP64): catches ctypes/C drift
    assert C.sizeof(L.StitchProfile) == 16
    assert C.sizeof(L.StitchImageDesc) == 32
    assert C.sizeof(L.StitchParams) == 32
    assert C.sizeof(L.StitchRect) == 32
    assert C.sizeof(L.StitchLayout) == 8 + 11 * 8 + 9 * 32 + 2 * 36
    assert C.sizeof(L.StitchImage) == 32 + 8 + 8 + 8 + 8
    assert C.sizeof(L.StitchJob) == 8 + 8 + 32 + 8 + 8 + 8 + 8 + 8 + 8 + 8 + 8 + 8 + 16
    assert C.sizeof(L.StitchImageInfo) == 16
    assert C.sizeof(L.StitchPreviewLayout) == 16 + 4 * 8
    assert C.sizeof(L.StitchResult) == C.sizeof(L.StitchLayout) + 8 + 16 + 8 + 8 + 8


def test_no_cpu_fallback_without_gpu():
    """On a box without a GPU the product refuses to run rather than falling back."""
    import torch
    if torch.cuda.is_available():
        return
    from imagestitching_b200 import StitchContext, StitchError
    try:
        StitchContext(0)
    except StitchError as e:
        assert e.status == L.E_CUDA
    else:
        raise AssertionError("StitchContext must fail without a CUDA device")


def test_product_does_not_import_oracle():
    """Nothing under imagestitching_b200/ may reference oracle/."""
    pkg = os.path.join(ROOT, "imagestitching_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".ts", ".cc")):
                src = open(os.path.join(dirpath, f), encoding="utf-8").read()
                for pat in (r"^\s*(from|import)\s+oracle\b", r"liboracle", r"oracle/", r"oracle_render"):
                    assert not re.search(pat, src, flags=re.M), (dirpath, f, pat)

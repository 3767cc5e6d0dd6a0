"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/coloc_b200.h declares; the product never touches oracle/; headers are plain C."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HDR = os.path.join(ROOT, "include", "coloc_b200.h")


def _declared():
    txt = open(HDR).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(clb_[a-z0-9_]+)\s*\(", txt)))


def test_header_is_plain_c(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "coloc_b200.h"\n#include "coloc_b200_latch_table.h"\n'
                   "int main(void){ clb_keypoint k; return sizeof(k) == 20 && sizeof(clb_partial) == 16 ? 0 : 1; }\n")
    exe = tmp_path / "t"
    subprocess.check_call(["/usr/bin/gcc", "-std=c99", "-Wall", "-Werror", "-Wno-unused-const-variable",
                           "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    assert subprocess.call([str(exe)]) == 0


def test_library_exports_every_declared_symbol():
    import coloc_b200
    p = coloc_b200.lib_path()
    if not os.path.exists(p):
        coloc_b200.build()
    lib = ctypes.CDLL(p)
    names = _declared()
    assert len(names) >= 25
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    assert lib.clb_version() >= 100


def test_errors_without_gpu_are_loud():
    """no CPU fallback: without a device the constructors fail with CLB_ECUDA, they do not degrade."""
    import torch
    import coloc_b200
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(coloc_b200.ClbError) as e:
        coloc_b200.Detector(640, 480)
    assert e.value.code == 2
    with pytest.raises(coloc_b200.ClbError):
        coloc_b200.Matcher()


def test_product_never_uses_oracle():
    bad = []
    for dp, _, fs in os.walk(os.path.join(ROOT, "coloc_b200")):
        if "build" in dp:
            continue
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                t = open(os.path.join(dp, f), errors="ignore").read()
                if re.search(r"import\s+oracle|from\s+oracle|coloc_oracle\.h|liboracle|oracle/", t):
                    bad.append(os.path.join(dp, f))
    assert not bad, bad


def test_keypoint_layout_matches_reference():
    """include/coloc/Keypoint.h:155-164 -> 20 bytes, x@0 y@4 score@8 angle@12 scale@16."""
    import coloc_b200
    dt = coloc_b200.KP_DTYPE
    assert dt.itemsize == 20
    assert [dt.fields[n][1] for n in ("x", "y", "score", "angle", "scale")] == [0, 4, 8, 12, 16]

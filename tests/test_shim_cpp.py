"""The drop-in C++ shim (coloc_b200/shim/coloc/GPUDetector.hpp, GPUMatcher.hpp): it must compile with the
reference's policy-template shells and link against the C ABI (CPU), and -- on the GPU box -- produce what the
oracle produces when driven the way coloc.hpp drives the reference classes."""
import os
import struct
import subprocess

import numpy as np
import pytest

from coloc_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "shim_driver.cpp")


def _build(out):
    import coloc_b200
    if not os.path.exists(coloc_b200.lib_path()):
        coloc_b200.build()
    libdir = os.path.join(ROOT, "coloc_b200")
    subprocess.check_call(["/usr/bin/g++", "-std=c++14", "-Wall", "-Werror", "-DCLB_SHIM_STANDALONE",
                           "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "coloc_b200", "shim"),
                           SRC, "-L", libdir, "-lcoloc_b200", f"-Wl,-rpath,{libdir}", "-o", out])


def test_shim_compiles_against_real_openmvg_headers(tmp_path):
    """the same shim + driver against the reference's OWN openMVG boundary types (regions_factory.hpp, indMatch.hpp, types.hpp
    from the submodule, Eigen and cereal vendored there) instead of the stand-ins -- only where /root/reference exists"""
    ref = "/root/reference/lib/openMVG/src"
    if not os.path.isdir(ref):
        pytest.skip("no /root/reference on this box")
    import coloc_b200
    if not os.path.exists(coloc_b200.lib_path()):
        coloc_b200.build()
    libdir = os.path.join(ROOT, "coloc_b200")
    exe = str(tmp_path / "shim_driver_omvg")
    subprocess.check_call(["/usr/bin/g++", "-std=c++14", "-Wall", "-Wno-deprecated-declarations", "-Wno-unused-local-typedefs", "-DCLB_SHIM_OPENMVG",
                           "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "coloc_b200", "shim"),
                           "-I", ref, "-I", os.path.join(ref, "third_party", "eigen"), "-I", os.path.join(ref, "dependencies", "cereal", "include"),
                           SRC, os.path.join(ref, "openMVG", "features", "feature.cpp"),      # SIOPointFeature's out-of-line members
                           "-L", libdir, "-lcoloc_b200", f"-Wl,-rpath,{libdir}", "-o", exe])
    assert subprocess.call([exe]) == 2


def test_shim_compiles_and_links(tmp_path):
    exe = str(tmp_path / "shim_driver")
    _build(exe)
    assert subprocess.call([exe]) == 2          # usage: no GPU call made


def _read_vec(buf, off, dtype):
    n = struct.unpack_from("<Q", buf, off)[0]
    off += 8
    a = np.frombuffer(buf, dtype=dtype, count=n, offset=off)
    return a, off + n * np.dtype(dtype).itemsize


@pytest.mark.gpu
def test_shim_matches_oracle(tmp_path, clb, oracle):
    w, h = 640, 480
    a, b, _ = synth.frame_pair(w, h, 2)
    paths = []
    for name, img in (("a", a), ("b", b)):
        p = str(tmp_path / f"{name}.pgm")
        with open(p, "wb") as f:
            f.write(b"P5\n%d %d\n255\n" % (w, h)); f.write(img.tobytes())
        paths.append(p)
    exe = str(tmp_path / "shim_driver"); out = str(tmp_path / "out.bin")
    _build(exe)
    subprocess.check_call([exe, paths[0], paths[1], str(w), str(h), out])
    buf = open(out, "rb").read()
    off = 0
    ka, off = _read_vec(buf, off, clb.KP_DTYPE); da, off = _read_vec(buf, off, np.uint64)
    kb, off = _read_vec(buf, off, clb.KP_DTYPE); db, off = _read_vec(buf, off, np.uint64)
    feats, off = _read_vec(buf, off, np.dtype([("x", "<f4"), ("y", "<f4"), ("s", "<f4"), ("o", "<f4")]))
    pair, off = _read_vec(buf, off, np.dtype([("i", "<u4"), ("j", "<u4")]))
    ratm, off = _read_vec(buf, off, np.dtype([("i", "<u4"), ("j", "<u4")]))
    mapm, off = _read_vec(buf, off, np.dtype([("i", "<u4"), ("j", "<u4")]))
    oka, oda = oracle.detect_describe(a); okb, odb = oracle.detect_describe(b)
    for k, ok in ((ka, oka), (kb, okb)):
        assert len(k) == len(ok)
        for fld in ("x", "y", "score", "angle", "scale"):
            assert np.array_equal(k[fld], ok[fld]), fld
    assert np.array_equal(da.reshape(-1, 8), oda) and np.array_equal(db.reshape(-1, 8), odb)
    # SIOPointFeature packing of GPUDetector.hpp:171-178
    sc = np.float32(1.2) ** oka["scale"].astype(np.float32)
    assert np.allclose(feats["x"], sc * oka["x"], rtol=1e-6) and np.allclose(feats["s"], 7 * sc, rtol=1e-6)
    assert np.array_equal(feats["o"], oka["angle"])
    # pair matches: (query=frame 0, train=frame 1), threshold 40, ascending query index
    m = oracle.match_2nn(oda, odb, 40)[0]
    q = np.nonzero(m >= 0)[0]
    assert np.array_equal(pair["i"], q) and np.array_equal(pair["j"], m[q])
    # map matches: map = frame 0, query = frame 1, threshold opts.thresh = 60, IndMatch(map idx, frame idx)
    mm = oracle.match_2nn(odb, oda, 60)[0]
    q = np.nonzero(mm >= 0)[0]
    assert np.array_equal(mapm["i"], mm[q]) and np.array_equal(mapm["j"], q)
    # CPU-build semantics: the reference's own DistanceRatioMatch on the same regions (when oracle/_ref travelled here)
    if oracle.ref_cpu() is not None and hasattr(oracle.ref_cpu(), "ref_distance_ratio_match"):
        from coloc_b200 import regions as R
        ref = oracle.ref_distance_ratio_match(oda, R.pack_regions(oka), odb, R.pack_regions(okb), 0.8)
        assert np.array_equal(np.stack([ratm["i"], ratm["j"]], 1), ref)

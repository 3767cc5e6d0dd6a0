"""Host side of the path (csrc/host_regions.cu): region packing, .feat/.desc files, DistanceRatioMatch's dedup tail.
CPU tests: no kernel runs.  Pinned against the reference's own compiled code (oracle/_ref) where it is present, and against
fixtures that code wrote (tests/golden/regions_*.{feat,desc,npz}, tools/gen_golden.py) everywhere."""
import os

import numpy as np
import pytest

import oracle as O
from coloc_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
needs_ref = pytest.mark.skipif(O.ref_cpu() is None or not hasattr(O.ref_cpu(), "ref_distance_ratio_match"), reason="oracle/_ref not built")


@pytest.fixture(scope="module")
def regions():
    from coloc_b200 import regions as R
    return R


def _kps(n, seed):
    from coloc_b200 import KP_DTYPE
    rng = np.random.default_rng(seed)
    k = np.zeros(n, KP_DTYPE)
    k["x"] = rng.integers(3, 1277, n); k["y"] = rng.integers(3, 717, n); k["score"] = rng.integers(41, 255, n)
    k["angle"] = rng.uniform(-np.pi, np.pi, n).astype(np.float32); k["scale"] = rng.integers(0, 8, n)
    return k


def test_pack_regions_formula(regions):
    k = _kps(500, 1)
    f = regions.pack_regions(k)
    s = np.power(np.float64(np.float32(1.2)), k["scale"].astype(np.float64)).astype(np.float32)     # (float)pow(1.2f, scale) in double
    assert np.array_equal(f[:, 0], s * k["x"].astype(np.float32)) and np.array_equal(f[:, 1], s * k["y"].astype(np.float32))
    assert np.array_equal(f[:, 2], np.float32(7.0) * s) and np.array_equal(f[:, 3].view(np.uint32), k["angle"].view(np.uint32))
    assert regions.pack_regions(k[:0]).shape == (0, 4)


def test_regions_files_match_golden(regions, tmp_path):
    g = np.load(os.path.join(GOLD, "regions_small.npz"))
    fp, dp = tmp_path / "a.feat", tmp_path / "a.desc"
    regions.save(fp, dp, g["feats"], g["desc"])
    assert open(fp, "rb").read() == open(os.path.join(GOLD, "regions_small.feat"), "rb").read()
    assert open(dp, "rb").read() == open(os.path.join(GOLD, "regions_small.desc"), "rb").read()
    f, d = regions.load(os.path.join(GOLD, "regions_small.feat"), os.path.join(GOLD, "regions_small.desc"))
    assert np.array_equal(d, g["desc"]) and np.array_equal(f.view(np.uint32), g["feats_loaded"].view(np.uint32))
    assert regions.count(os.path.join(GOLD, "regions_small.desc")) == len(g["desc"])


@needs_ref
def test_regions_files_equal_reference_writer(regions, tmp_path):
    k = _kps(777, 2)
    feats = regions.pack_regions(k); feats[5] = (1e-7, 123456.789, 3.0e10, -0.0)       # formatting corner cases
    desc = synth.random_descriptors(777, 3)
    a = (tmp_path / "a.feat", tmp_path / "a.desc"); b = (tmp_path / "b.feat", tmp_path / "b.desc")
    regions.save(a[0], a[1], feats, desc)
    assert O.ref_regions_save(b[0], b[1], feats, desc)
    assert open(a[0], "rb").read() == open(b[0], "rb").read() and open(a[1], "rb").read() == open(b[1], "rb").read()
    f1, d1 = regions.load(b[0], b[1]); f2, d2 = O.ref_regions_load(a[0], a[1], 1000)
    assert np.array_equal(d1, desc) and np.array_equal(d2, desc) and np.array_equal(f1.view(np.uint32), f2.view(np.uint32))
    # empty regions
    regions.save(a[0], a[1], feats[:0], desc[:0]); assert O.ref_regions_save(b[0], b[1], feats[:0], desc[:0])
    assert open(a[1], "rb").read() == open(b[1], "rb").read() == (0).to_bytes(8, "little")
    f, d = regions.load(a[0], a[1]); assert len(f) == 0 and len(d) == 0


def test_regions_load_errors(regions, tmp_path):
    from coloc_b200 import ClbError
    with pytest.raises(ClbError) as e:
        regions.load(None, tmp_path / "missing.desc")
    assert e.value.code == 7
    feats = np.zeros((3, 4), np.float32); desc = synth.random_descriptors(3, 1)
    regions.save(tmp_path / "x.feat", tmp_path / "x.desc", feats, desc)
    regions.save(None, tmp_path / "y.desc", None, desc[:2])
    with pytest.raises(ClbError):
        regions.load(tmp_path / "x.feat", tmp_path / "y.desc")          # 3 features, 2 descriptors


def _dedup_case(seed, nI=600, nJ=500, gx=12, gy=9):
    """regions with many repeated (x, y) positions (pyramid levels re-detect the same corner) and repeated descriptor rows"""
    rng = np.random.default_rng(seed)
    dI = synth.random_descriptors(nI, seed + 1)
    dJ, _ = synth.planted_queries(dI, nJ, seed + 2)
    dJ[nJ // 2:nJ // 2 + 40] = dJ[:40]                                  # repeated queries -> several j for one i
    fI = np.zeros((nI, 4), np.float32); fJ = np.zeros((nJ, 4), np.float32)
    fI[:, 0] = rng.integers(0, gx, nI) * 1.5; fI[:, 1] = rng.integers(0, gy, nI) * 2.25
    fJ[:, 0] = rng.integers(0, gx, nJ) * 1.5; fJ[:, 1] = rng.integers(0, gy, nJ) * 2.25
    fJ[nJ // 2:nJ // 2 + 40, :2] = fJ[:40, :2]
    return dI, fI, dJ, fJ


def test_dedup_matches_golden(regions):
    g = np.load(os.path.join(GOLD, "regions_dedup.npz"))
    got = regions.dedup_matches(g["match"], g["feats_i"], g["feats_j"])
    assert np.array_equal(got, g["pairs"])
    assert len(got) < int((g["match"] >= 0).sum())                      # the fixture does exercise the (x, y) dedup
    assert regions.dedup_matches(np.full(10, -1, np.int32), g["feats_i"], g["feats_j"]).shape == (0, 2)


@needs_ref
@pytest.mark.parametrize("seed,gx,gy", [(1, 12, 9), (2, 300, 200), (3, 40, 2000), (4, 2000, 30), (5, 100000, 100000)])
def test_dedup_equals_reference(regions, seed, gx, gy):
    dI, fI, dJ, fJ = _dedup_case(seed, gx=gx, gy=gy)
    idx, dist, ok = O.ref_bruteforce_2nn(dI, dJ, 0.8)
    match = np.where(ok.astype(bool), idx[:, 0], -1).astype(np.int32)
    got = regions.dedup_matches(match, fI, fJ)
    ref = O.ref_distance_ratio_match(dI, fI, dJ, fJ, 0.8)
    assert np.array_equal(got, ref)
    assert len(ref) > 0


@needs_ref
@pytest.mark.gpu
@pytest.mark.parametrize("seed,nI,nJ", [(4, 600, 500), (5, 3000, 2500), (6, 2, 5), (7, 1, 5)])
def test_distance_ratio_match_equals_reference(clb, regions, seed, nI, nJ):
    dI, fI, dJ, fJ = _dedup_case(seed, nI, max(nJ, 80), gx=500, gy=400) if nI >= 80 else (synth.random_descriptors(nI, seed), np.zeros((nI, 4), np.float32),
                                                                          synth.random_descriptors(nJ, seed + 1), np.ones((nJ, 4), np.float32))
    m = clb.Matcher(maxkp=4096)
    got = regions.distance_ratio_match(m, dI, fI, dJ, fJ, 0.8)
    ref = O.ref_distance_ratio_match(dI, fI, dJ, fJ, 0.8)
    assert np.array_equal(got, ref)
    m.close()

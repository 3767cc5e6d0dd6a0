#!/usr/bin/env python3
"""Generate tests/golden/*.npz.  Run in the build container (where /root/reference exists): the
keypoints come from the REFERENCE's compiled KFAST.h + FeatureAngle.h (oracle/_ref) and are asserted
equal to the restatement before being written; descriptors / pyramid / matches come from the
restatement (the reference's CUDA kernels cannot run here: no GPU).  The fixtures then pin the
oracle on machines without /root/reference, and the CUDA path through tests/test_gpu_parity.py."""
import hashlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle as O
from coloc_b200 import synth

GOLD = os.path.join(ROOT, "tests", "golden")
os.makedirs(GOLD, exist_ok=True)
assert O.ref_cpu() is not None, "build oracle/_ref first (make -C oracle ref_cpu)"



def gen_frame(name, w, h, seed, thresh, need_quirk):
    """one small frame through the whole detect + describe restatement, every level's keypoints + angles cross-checked against the
    reference's compiled CPU code before anything is written"""
    img = synth.textured_frame(w, h, seed)
    kps, desc, pyr = O.detect_describe(img, thresh=thresh, want_pyramid=True)
    bites = False
    for L, lv in enumerate(pyr):
        ref = O.ref_kfast(lv, thresh)
        mine = kps[kps["scale"] == L]
        assert len(ref) == len(mine) and np.array_equal(ref["x"], mine["x"]) and np.array_equal(ref["y"], mine["y"]) \
            and np.array_equal(ref["score"], mine["score"]), L
        for k in mine:
            assert np.float32(O.ref_feature_angle(lv, k["x"], k["y"])).view(np.uint32) == k["angle"].view(np.uint32)
        bites = bites or not np.array_equal(O.fast_nms(lv, thresh, quirk=0), O.fast_nms(lv, thresh, quirk=1))
    assert bites == need_quirk, "the frame must (not) have a level on which KFAST.h:243 drops columns"
    np.savez_compressed(os.path.join(GOLD, name), img=img, thresh=thresh, kps=kps, desc=desc,
                        pyr_sha256=hashlib.sha256(b"".join(p.tobytes() for p in pyr)).hexdigest())
    print(name, len(kps), "keypoints; levels", [lv.shape[1] for lv in pyr])


if "frame214" in sys.argv[1:]:          # only the fixture added in round 2: level 0 is 214 px wide, (214 - 38) % 16 == 0
    gen_frame("frame214.npz", 214, 161, 78, 20, True)
    sys.exit(0)

img = synth.textured_frame(168, 126, 77)
thresh = 20
kps, desc, pyr = O.detect_describe(img, thresh=thresh, want_pyramid=True)
# cross-check every level's keypoints + angles against the reference's compiled CPU code
off = 0
for L, lv in enumerate(pyr):
    ref = O.ref_kfast(lv, thresh)
    mine = kps[kps["scale"] == L]
    assert len(ref) == len(mine) and np.array_equal(ref["x"], mine["x"]) and np.array_equal(ref["y"], mine["y"]) \
        and np.array_equal(ref["score"], mine["score"]), L
    for k in mine:
        assert np.float32(O.ref_feature_angle(lv, k["x"], k["y"])).view(np.uint32) == k["angle"].view(np.uint32)
np.savez_compressed(os.path.join(GOLD, "frame168.npz"), img=img, thresh=thresh, kps=kps, desc=desc,
                    pyr_sha256=hashlib.sha256(b"".join(p.tobytes() for p in pyr)).hexdigest())
print("frame168:", len(kps), "keypoints")

T = synth.random_descriptors(257, 21)
Q, src = synth.planted_queries(T, 96, 22)
Q[:8] = T[:8]; T[100] = T[3]          # exact hits + a duplicated row (tie on best)
m, b, s = O.match_2nn(Q, T, 40)
idx, dist, ok = O.ref_bruteforce_2nn(T, Q, 0.8)
assert np.array_equal(dist[:, 0], b) and np.array_equal(dist[:, 1], s)
np.savez_compressed(os.path.join(GOLD, "match_small.npz"), Q=Q, T=T, threshold=40, match=m, best=b, second=s,
                    ratio_ok=ok)
print("match_small:", int((m >= 0).sum()), "matches of", len(Q))

# ---- host side: .feat/.desc written by the REFERENCE's Binary_Regions::Save, and DistanceRatioMatch's dedup tail
from coloc_b200 import regions as R, KP_DTYPE
rng = np.random.default_rng(5)
k = np.zeros(64, KP_DTYPE)
k["x"] = rng.integers(3, 637, 64); k["y"] = rng.integers(3, 477, 64); k["scale"] = rng.integers(0, 8, 64)
k["angle"] = rng.uniform(-np.pi, np.pi, 64).astype(np.float32)
feats = R.pack_regions(k); feats[3] = (1e-7, 123456.789, 3.0e10, -0.0)
desc = synth.random_descriptors(64, 6)
fp, dp = os.path.join(GOLD, "regions_small.feat"), os.path.join(GOLD, "regions_small.desc")
assert O.ref_regions_save(fp, dp, feats, desc)
fl, dl = O.ref_regions_load(fp, dp, 100)
assert np.array_equal(dl, desc)
np.savez_compressed(os.path.join(GOLD, "regions_small.npz"), feats=feats, desc=desc, feats_loaded=fl)

sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_host_regions import _dedup_case
dI, fI, dJ, fJ = _dedup_case(9, gx=60, gy=45)
idx, dist, ok = O.ref_bruteforce_2nn(dI, dJ, 0.8)
match = np.where(ok.astype(bool), idx[:, 0], -1).astype(np.int32)
pairs = O.ref_distance_ratio_match(dI, fI, dJ, fJ, 0.8)
np.savez_compressed(os.path.join(GOLD, "regions_dedup.npz"), match=match, feats_i=fI, feats_j=fJ, pairs=pairs)
print("regions_dedup:", int((match >= 0).sum()), "ratio matches ->", len(pairs), "after dedup")

# ---- two-view residuals from the REFERENCE's own functors (oracle/_ref), 4 error models
from test_geometry import scene, hypotheses
x1, x2, Hm, rng = scene(400, 3)
out = {"x1": x1, "x2": x2}
for kind in range(4):
    Ms = hypotheses(kind, Hm, rng, 4)
    out[f"models{kind}"] = Ms
    out[f"res{kind}"] = np.stack([O.ref_model_residuals(kind, M, x1, x2) for M in Ms])
np.savez_compressed(os.path.join(GOLD, "residuals.npz"), **out)
print("residuals:", {k: v.shape for k, v in out.items() if k.startswith("res")})

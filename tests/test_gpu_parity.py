"""GPU parity tests (the gate): the CUDA path, called through the C ABI, against the CPU oracle on
identical seeded inputs, against the committed golden fixtures, and -- at BASELINE sizes --
through size-independent properties.  Integer/byte/index work is compared bit-exactly; the only
floating-point outputs (angle, pyramid) are also required to be bit-identical to the oracle's
explicit-fma restatement (tolerance stated per test; north_star allows 1 LSB on pyramid pixels)."""
import os

import numpy as np
import pytest

from coloc_b200 import synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="module")
def matcher(clb):
    m = clb.Matcher(maxkp=4096)
    yield m
    m.close()


def _check_match(clb, oracle, matcher, Q, T, thr):
    got = matcher.compute_matches(Q, T, thr, want_dist=True)
    ref = oracle.match_2nn(Q, T, thr)
    for name, a, b in zip(("match", "best", "second"), got, ref):
        assert np.array_equal(a, b), (name, int((a != b).sum()))


# ------------------------------------------------------------------ Hamming 2-NN
@pytest.mark.parametrize("nq,nt", [(1, 1), (1, 2), (3, 1), (7, 5), (33, 64), (64, 65), (255, 257), (300, 1000),
                                   (1000, 777), (2049, 4100), (5000, 5000)])
@pytest.mark.parametrize("hs", [2, 0, 1, 3])
def test_match_vs_oracle(clb, oracle, matcher, nq, nt, hs):
    if hs != 2 and nq > 1000:
        pytest.skip("knob sweep on small sizes only")
    if hs != 2 and not clb.is_lab():
        pytest.skip("carry-save depth variants exist in the lab build only (CLB_LIB=lab)")
    if clb.is_lab():
        clb.set_param("match_hs", hs)
    try:
        T = synth.random_descriptors(nt, 100 + nt)
        Q, _ = synth.planted_queries(T, nq, 200 + nq)
        rng = np.random.default_rng(nq * 7 + nt)
        k = max(1, nq // 4)
        Q[:k] = T[rng.integers(0, nt, k)]                    # exact hits
        if nt > 4:
            T[nt // 2] = T[1]; T[nt - 1] = T[1]              # duplicated rows -> ties, first index must win
        for thr in (40, 60, 0):
            _check_match(clb, oracle, matcher, Q, T, thr)
    finally:
        if clb.is_lab():
            clb.set_param("match_hs", 2)


@pytest.mark.parametrize("rq", [1, 2, 4])
@pytest.mark.parametrize("chunks", [1, 3, 0])
def test_match_knobs(clb, oracle, matcher, rq, chunks):
    clb.set_param("match_rq", rq); clb.set_param("match_chunks", chunks)
    try:
        T = synth.random_descriptors(1500, 1); Q, _ = synth.planted_queries(T, 700, 2)
        T[900] = T[10]
        _check_match(clb, oracle, matcher, Q, T, 40)
    finally:
        clb.set_param("match_rq", 0); clb.set_param("match_chunks", 0)


def test_match_edge_cases(clb, oracle, matcher):
    T = synth.random_descriptors(8, 5)
    # empty DB: no match, sentinels of CUDAK2NN.cu:54
    m, b, s = matcher.compute_matches(T[:3], T[:0], 40, want_dist=True)
    assert list(m) == [-1] * 3 and list(b) == [100000] * 3 and list(s) == [200000] * 3
    # empty query set
    assert len(matcher.compute_matches(T[:0], T, 40)) == 0
    # one DB row: always matched
    m, b, s = matcher.compute_matches(T[:4], T[:1], 40, want_dist=True)
    assert list(m) == [0] * 4 and list(s) == [100000] * 4 and b[0] == 0
    # all-zero and all-one rows: distance 512
    Z = np.zeros((1, 8), np.uint64); F = np.full((2, 8), np.uint64(0xFFFFFFFFFFFFFFFF))
    m, b, s = matcher.compute_matches(Z, F, 40, want_dist=True)
    assert b[0] == 512 and s[0] == 512 and m[0] == -1
    _check_match(clb, oracle, matcher, T, T, 40)             # self match: distance 0 everywhere


def test_match_golden(clb, matcher):
    g = np.load(os.path.join(GOLD, "match_small.npz"))
    m, b, s = matcher.compute_matches(g["Q"], g["T"], int(g["threshold"]), want_dist=True)
    assert np.array_equal(m, g["match"]) and np.array_equal(b, g["best"]) and np.array_equal(s, g["second"])
    # ratio mode reproduces openMVG's NNdistanceRatio decisions (generated from the reference's compiled code)
    mr = matcher.compute_matches_ratio(g["Q"], g["T"], 0.8)
    assert np.array_equal(mr >= 0, g["ratio_ok"].astype(bool))


def test_match_map_api(clb, oracle, matcher):
    M = synth.random_descriptors(3000, 9); Q, src = synth.planted_queries(M, 500, 10)
    matcher.set_map_data(M)
    got = matcher.match_with_map(Q, want_dist=True)          # threshold from opts (60), GPUMatcher.hpp:256
    ref = oracle.match_2nn(Q, M, 60)
    assert all(np.array_equal(a, b) for a, b in zip(got, ref))
    t = src >= 0
    assert (got[0][t] == src[t]).mean() > 0.95               # planted matches recovered


def test_match_full_size_properties(clb, oracle):
    """C4-shaped run (50k queries; DB cut to 400k rows to keep the test short): properties that do not need the
    oracle at full size + an oracle check on a 64-query slice."""
    nq, nt = 50000, 400000
    T = synth.random_descriptors(nt, 42); Q, src = synth.planted_queries(T, nq, 43)
    m = clb.Matcher(maxkp=nt)
    match, best, second = m.compute_matches(Q, T, 40, want_dist=True)
    t = src >= 0
    assert (match[t] == src[t]).mean() > 0.999               # d~51 vs random ~256: all planted rows found
    assert (match[~t] == -1).mean() > 0.99
    assert (best <= second).all() and (best[t] < 120).all()
    # linearity: XOR-ing every row with the same mask leaves all distances unchanged
    mask = synth.random_descriptors(1, 7)
    m2, b2, s2 = m.compute_matches(Q[:5000] ^ mask, T ^ mask, 40, want_dist=True)
    assert np.array_equal(m2, match[:5000]) and np.array_equal(b2, best[:5000]) and np.array_equal(s2, second[:5000])
    # oracle on a slice
    sl = slice(1000, 1064)
    ro = oracle.match_2nn(Q[sl], T, 40)
    assert np.array_equal(ro[0], match[sl]) and np.array_equal(ro[1], best[sl]) and np.array_equal(ro[2], second[sl])
    m.close()


def test_partial_and_merge(clb, oracle):
    """DB-row sharding on one GPU: partials of 3 uneven shards + merge == single scan."""
    import torch
    nq, nt = 1200, 5000
    T = synth.random_descriptors(nt, 3); Q, _ = synth.planted_queries(T, nq, 4)
    T[4000] = T[10]; T[2500] = T[10]                          # ties across shards: lowest global index wins
    Q[0] = T[10]
    m = clb.Matcher(maxkp=nt)
    dQ = torch.from_numpy(Q.view(np.int64)).cuda(); dT = torch.from_numpy(T.view(np.int64)).cuda()
    bounds = [0, 1700, 1700, 3301, nt]                        # includes an empty shard
    parts = torch.zeros((len(bounds) - 1, nq, 4), dtype=torch.int32, device="cuda")
    for i in range(len(bounds) - 1):
        a, b = bounds[i], bounds[i + 1]
        m.match_partial_dev(dQ, nq, dT[a:] if b > a else dT, b - a, a, parts[i])
    out = torch.zeros((3, nq), dtype=torch.int32, device="cuda")
    m.merge_partials_dev(parts, len(bounds) - 1, nq, 40, out[0], out[1], out[2])
    torch.cuda.synchronize()
    ref = oracle.match_2nn(Q, T, 40)
    got = out.cpu().numpy()
    for i in range(3):
        assert np.array_equal(got[i], ref[i]), i
    m.close()


# ------------------------------------------------------------------ tensor-core path (match_tc.cu)
@pytest.fixture()
def tc_forced(clb):
    """force the tcgen05 kernels at every size (the default picks them from 2^24 comparisons up)"""
    clb.set_param("match_tc", 1)
    yield
    clb.set_param("match_tc", -1); clb.set_param("match_tc_chunks", 0)
    if clb.is_lab():
        clb.set_param("match_tc_ver", 4)


@pytest.mark.parametrize("ver", [4, 3, 2, 1])
@pytest.mark.parametrize("nq,nt", [(1, 1), (1, 2), (3, 1), (7, 5), (33, 64), (64, 65), (128, 128), (255, 257), (256, 129),
                                   (300, 1000), (1000, 777), (2049, 4100), (5000, 5000)])
def test_match_tc_vs_oracle(clb, oracle, matcher, tc_forced, nq, nt, ver):
    if ver != 4 and not clb.is_lab():
        pytest.skip("superseded tensor-core kernels exist in the lab build only (CLB_LIB=lab)")
    if clb.is_lab():
        clb.set_param("match_tc_ver", ver)
    T = synth.random_descriptors(nt, 100 + nt)
    Q, _ = synth.planted_queries(T, nq, 200 + nq)
    rng = np.random.default_rng(nq * 7 + nt)
    k = max(1, nq // 4)
    Q[:k] = T[rng.integers(0, nt, k)]                        # exact hits (distance 0)
    if nt > 4:
        T[nt // 2] = T[1]; T[nt - 1] = T[1]                  # duplicated rows -> ties, first index must win
    for thr in (40, 60, 0):
        _check_match(clb, oracle, matcher, Q, T, thr)
    assert clb.get_param("last_match_path") == ver


def test_match_tc_ragged_tail_race(clb, oracle, matcher, tc_forced):
    """20 000 x 20 000: 157 DB tiles (ragged last tile) over several chunks; the shape that exposed the v4 staging race"""
    T = synth.random_descriptors(20000, 31); Q, _ = synth.planted_queries(T, 20000, 32)
    for _ in range(3):
        got = matcher.compute_matches(Q, T, 40, want_dist=True)
        clb.set_param("match_tc", 0)
        ref = matcher.compute_matches(Q, T, 40, want_dist=True)
        clb.set_param("match_tc", 1)
        assert all(np.array_equal(a, b) for a, b in zip(got, ref))
    sl = slice(16000, 16064)
    ro = oracle.match_2nn(Q[sl], T, 40)
    assert all(np.array_equal(a[sl], b) for a, b in zip(got, ro))


@pytest.mark.parametrize("chunks", [1, 2, 5])
def test_match_tc_chunking(clb, oracle, matcher, tc_forced, chunks):
    """DB cut into 1 / 2 / 5 chunks over blockIdx.y: merge keeps first-min-wins across chunk borders"""
    clb.set_param("match_tc_chunks", chunks)
    T = synth.random_descriptors(3000, 11); Q, _ = synth.planted_queries(T, 700, 12)
    T[2900] = T[10]; T[1500] = T[10]; Q[0] = T[10]
    _check_match(clb, oracle, matcher, Q, T, 40)


def test_match_tc_edge_cases(clb, oracle, matcher, tc_forced):
    test_match_edge_cases(clb, oracle, matcher)
    # adversarial order: every DB row is closer than the one before it (the epilogue's skip test never fires)
    nt = 2048
    q = synth.random_descriptors(1, 77)
    bits = np.unpackbits(q.view(np.uint8), bitorder="little").reshape(1, 512)
    Tb = np.repeat(bits, nt, axis=0)
    for r in range(nt):
        Tb[r, :511 - (r * 511) // nt] ^= 1
    T = np.packbits(Tb, axis=1, bitorder="little").view(np.uint64)
    _check_match(clb, oracle, matcher, np.repeat(q, 130, axis=0), T, 0)


def test_match_tc_golden_map_and_shards(clb, oracle, matcher, tc_forced):
    test_match_golden(clb, matcher)
    test_match_map_api(clb, oracle, matcher)
    # a second query set against the same map reuses the cached expansion; a new map invalidates it
    M2 = synth.random_descriptors(2000, 19); Q2, _ = synth.planted_queries(M2, 300, 20)
    matcher.set_map_data(M2)
    for Qs in (Q2, Q2[::-1].copy()):
        got = matcher.match_with_map(Qs, 60, want_dist=True)
        ref = oracle.match_2nn(Qs, M2, 60)
        assert all(np.array_equal(a, b) for a, b in zip(got, ref))
    test_partial_and_merge(clb, oracle)


def test_match_default_path_is_tensor_core_at_size(clb, matcher):
    T = synth.random_descriptors(5000, 1); Q = synth.random_descriptors(5000, 2)
    matcher.compute_matches(Q, T, 40)
    assert clb.get_param("last_match_path") == 4              # 2.5e7 comparisons >= 2^24
    matcher.compute_matches(Q[:100], T[:100], 40)
    assert clb.get_param("last_match_path") == 0


# ------------------------------------------------------------------ pyramid
@pytest.mark.parametrize("w,h", [(640, 480), (1280, 720), (333, 251), (64, 48)])
def test_pyramid_vs_oracle(clb, oracle, w, h):
    img = synth.textured_frame(w, h, 5)
    det = clb.Detector(w, h, maxkp=1000)
    got = det.pyramid(img)
    ref = oracle.pyramid(img)
    for L, (a, b) in enumerate(zip(got, ref)):
        assert a.shape == b.shape
        # north_star tolerance is 1 LSB; the explicit-fma tree is in fact bit-identical
        assert np.array_equal(a, b), (L, int(np.abs(a.astype(int) - b.astype(int)).max()))
    det.close()


# ------------------------------------------------------------------ FAST + NMS + angle
FAST_VARS = [55, 23, 19, 7, 3, 1, 2, 0]     # fast_var (lab build): queued score kernel / fused tail / 64x32 tiles / 16-bit-lane pre-test / first generation


def _fast_vars(clb):
    return FAST_VARS if clb.is_lab() else [55]


def _set_fast_var(clb, var):
    if clb.is_lab():
        clb.set_param("fast_var", var)


# widths 38, 54, 70, 214, 1286: (w - 38) % 16 == 0, the levels on which the compiled reference's `1 << 32` (KFAST.h:243) bites
@pytest.mark.parametrize("w,h,thr", [(640, 480, 40), (533, 400, 40), (1067, 600, 40), (257, 193, 20), (70, 50, 10),
                                     (214, 161, 20), (214, 161, 40), (1286, 723, 40), (38, 30, 5), (54, 40, 8), (39, 20, 5),
                                     (9, 9, 5), (7, 7, 1), (6, 30, 5)])
@pytest.mark.parametrize("quirk", [1, 0])
def test_fast_vs_oracle(clb, oracle, w, h, thr, quirk):
    img = synth.textured_frame(w, h, 3)
    det = clb.Detector(max(w, 8), max(h, 8), maxkp=w * h // 4 + 16, thresh=thr, fast_quirk=quirk)
    ref = oracle.fast_nms(img, thr, quirk=quirk)
    for k in ref:
        k["angle"] = oracle.feature_angle(img, k["x"], k["y"]); k["scale"] = 2
    try:
        for var in _fast_vars(clb):
            _set_fast_var(clb, var)
            got = det.fast_nms_angle(img, level_idx=2)
            assert len(got) == len(ref), (var, len(got), len(ref))
            assert np.array_equal(got.view(np.uint8), ref.view(np.uint8)), var   # x, y, score, angle bits, scale, order
    finally:
        _set_fast_var(clb, 55)
    det.close()


@pytest.mark.parametrize("w0,h0,k", [(640, 480, 0), (640, 480, 1), (3840, 2160, 0)])
def test_fast_equals_compiled_reference_on_quirk_levels(clb, oracle, w0, h0, k):
    """level 6 of the C1 / C5 frames (214 x 161, 1286 x 723): the CUDA path with fast_quirk = 1 (the default) against the
    reference's own KFAST.h compiled unmodified (oracle/_ref), not only against the restatement; and the two semantics differ
    there, so the test cannot pass by accident."""
    if oracle.ref_cpu() is None:
        pytest.skip("oracle/_ref/libcoloc_ref_cpu.so not built")
    img = synth.config_frame(w0, h0, k)
    lv = oracle.pyramid(img)[6]
    assert (lv.shape[1] - 38) % 16 == 0
    det = clb.Detector(lv.shape[1], lv.shape[0], maxkp=lv.size // 4 + 16, thresh=40)
    got = det.fast_nms_angle(lv, level_idx=6)
    ref = oracle.ref_kfast(lv, 40)
    assert len(got) == len(ref) > 100
    for f in ("x", "y", "score"):
        assert np.array_equal(got[f], ref[f]), f
    det.set_param("fast_quirk", 0)
    q0 = det.fast_nms_angle(lv, level_idx=6)
    assert len(q0) > len(got)                                    # the intended semantics keep the columns the reference drops
    o0 = oracle.fast_nms(lv, 40, quirk=0)
    assert len(q0) == len(o0) and all(np.array_equal(q0[f], o0[f]) for f in ("x", "y", "score"))
    det.close()


def test_fast_noise_and_flat(clb, oracle):
    # uniform noise: every position passes the pre-test, ~1/3 are corners -- the shared-memory queues of the v2 kernel run full
    det = clb.Detector(640, 480, maxkp=200000, thresh=40)
    noise = np.random.default_rng(0).integers(0, 256, (480, 640), dtype=np.uint8)
    ref = oracle.fast_nms(noise, 40)
    try:
        for var in ([55, 23, 7, 3, 0] if clb.is_lab() else [55]):
            _set_fast_var(clb, var)
            got = det.fast_nms_angle(noise)
            assert len(got) == len(ref) > 20000 and np.array_equal(got["x"], ref["x"]) and np.array_equal(got["score"], ref["score"])
            assert len(det.fast_nms_angle(np.full((480, 640), 128, np.uint8))) == 0
    finally:
        _set_fast_var(clb, 55)
    det.close()
    # noise on a quirk width: every 16-column block has pre-test passes, so the walk advances by 32 from column 3 and lands on
    # w - 35 exactly when (w - 38) % 32 == 0 -- every row of a 230-wide image is truncated, none of a 214-wide one
    for w in (230, 214):
        noise = np.random.default_rng(w).integers(0, 256, (60, w), dtype=np.uint8)
        det = clb.Detector(w, 60, maxkp=20000, thresh=40)
        got = det.fast_nms_angle(noise); ref = oracle.fast_nms(noise, 40, quirk=1)
        assert len(got) == len(ref) and np.array_equal(got["x"], ref["x"]) and np.array_equal(got["y"], ref["y"])
        assert (got["x"].max() < w - 35) == (w == 230)
        if oracle.ref_cpu() is not None:
            rk = oracle.ref_kfast(noise, 40)
            assert np.array_equal(got["x"], rk["x"]) and np.array_equal(got["y"], rk["y"]) and np.array_equal(got["score"], rk["score"])
        det.close()


def test_fast_threshold_zero_full_queues(clb, oracle):
    # thresh 1 on noise: the candidate queue holds (nearly) every position of every tile
    det = clb.Detector(200, 150, maxkp=40000, thresh=1)
    noise = np.random.default_rng(5).integers(0, 256, (150, 200), dtype=np.uint8)
    got = det.fast_nms_angle(noise)
    ref = oracle.fast_nms(noise, 1)
    assert len(got) == len(ref) and np.array_equal(got["x"], ref["x"]) and np.array_equal(got["y"], ref["y"]) \
        and np.array_equal(got["score"], ref["score"])
    det.close()


def test_fast_capacity_error(clb):
    det = clb.Detector(640, 480, maxkp=100, thresh=40)
    noise = np.random.default_rng(0).integers(0, 256, (480, 640), dtype=np.uint8)
    with pytest.raises(clb.ClbError) as e:
        det.fast_nms_angle(noise)
    assert e.value.code == 3
    det.close()


# ------------------------------------------------------------------ LATCH describe
def test_sincos_bit_exact(clb, oracle):
    det = clb.Detector(64, 64, maxkp=16)
    a = np.concatenate([np.linspace(-np.pi, np.pi, 400001), [0.0, np.pi, -np.pi, 1e-8, -1e-8]]).astype(np.float32)
    s, c = det.sincos(a)
    so, co = oracle.sincos_many(a)
    assert np.array_equal(s.view(np.uint32), so.view(np.uint32)) and np.array_equal(c.view(np.uint32), co.view(np.uint32))
    det.close()


@pytest.mark.parametrize("w,h", [(640, 480), (320, 240)])
def test_describe_vs_oracle(clb, oracle, w, h):
    img = synth.textured_frame(w, h, 8)
    pyr = oracle.pyramid(img)
    kps, _ = oracle.detect_describe(img, thresh=30, cap=50000)
    # add border keypoints (clamp-to-edge sampling) and arbitrary angles
    extra = np.zeros(6, kps.dtype)
    extra["x"] = [0, w - 1, 3, w - 4, 10, 5]; extra["y"] = [0, h - 1, 3, h - 4, 2, h - 2]
    extra["angle"] = [0.0, 3.1415927, -3.1415927, 0.7853982, -1.5707964, 2.5]
    kps = np.concatenate([kps, extra])
    det = clb.Detector(w, h, maxkp=len(kps) + 1)
    got = det.describe(pyr, kps)
    ref = oracle.latch_describe(pyr, kps)
    bad = (got != ref).any(axis=1)
    assert not bad.any(), (int(bad.sum()), len(kps))
    det.close()


# ------------------------------------------------------------------ whole frame
@pytest.mark.parametrize("w,h", [(640, 480), (1280, 720)])
def test_detect_describe_end_to_end(clb, oracle, w, h):
    a, b, H = synth.frame_pair(w, h, 0)
    det = clb.Detector(w, h, maxkp=100000)
    for img in (a, b):
        kps, desc = det.detect_and_describe(img)
        okps, odesc = oracle.detect_describe(img, cap=100000)          # quirk = 1: the compiled reference's keypoint set
        assert len(kps) == len(okps) > 1000
        assert np.array_equal(kps.view(np.uint8), okps.view(np.uint8))
        assert np.array_equal(desc, odesc)
    det.close()


def test_config1_frame_equals_reference_semantics(clb, oracle):
    """BASELINE config 1 frame (640x480, level 6 is 214 wide): 5 207 keypoints with the reference's column walk (the default),
    5 232 with every column scanned; both bit-exact against the oracle, keypoints and descriptors"""
    img = synth.config_frame(640, 480, 0)
    counts = {}
    for quirk in (1, 0):
        det = clb.Detector(640, 480, maxkp=50000, fast_quirk=quirk)
        kps, desc = det.detect_and_describe(img)
        okps, odesc = oracle.detect_describe(img, quirk=quirk, cap=50000)
        assert np.array_equal(kps.view(np.uint8), okps.view(np.uint8)) and np.array_equal(desc, odesc)
        counts[quirk] = len(kps)
        det.close()
    assert counts == {1: 5207, 0: 5232}, counts


def test_config5_frame_end_to_end(clb, oracle):
    # BASELINE config 5 frame size: 3840x2160 (60 tiles per row: two words of tile-occupancy bitmap per row, 32-bit level offsets
    # beyond 8 MB, the dense and the sparse compaction path in one frame) against the oracle, keypoints and descriptors bit for bit
    w, h = 3840, 2160
    img = synth.config_frame(w, h, 0)
    det = clb.Detector(w, h, maxkp=200000)
    kps, desc = det.detect_and_describe(img)
    okps, odesc = oracle.detect_describe(img, cap=200000)              # quirk = 1 (level 6 is 1286 wide)
    assert len(kps) == len(okps) == 50009
    assert np.array_equal(kps.view(np.uint8), okps.view(np.uint8))
    assert np.array_equal(desc, odesc)
    det.set_param("fast_quirk", 0)
    k0, d0 = det.detect_and_describe(img)
    ok0, od0 = oracle.detect_describe(img, quirk=0, cap=200000)
    assert len(k0) == len(ok0) == 50108 and np.array_equal(k0.view(np.uint8), ok0.view(np.uint8)) and np.array_equal(d0, od0)
    det.set_param("fast_quirk", 1)
    # a mostly flat frame with a few textured islands: nearly every row takes the sparse path
    flat = np.full((h, w), 90, np.uint8)
    for (y0, x0) in ((100, 64), (900, 2050), (1500, 3700), (2100, 0), (40, 3776)):
        flat[y0:y0 + 48, x0:x0 + 64] = img[y0:y0 + 48, x0:x0 + 64]
    kps, desc = det.detect_and_describe(flat)
    okps, odesc = oracle.detect_describe(flat, cap=200000)
    assert len(kps) == len(okps) > 20
    assert np.array_equal(kps.view(np.uint8), okps.view(np.uint8)) and np.array_equal(desc, odesc)
    det.close()


def test_alternating_frames_reuse_buffers(clb, oracle):
    """One detector, one geometry, DIFFERENT frame contents back to back (the pyramid, score map and keypoint buffers are
    reused with new contents; the frame's kernels are a programmatic-dependent-launch chain): every frame equals the oracle.
    Catches a consumer kernel that reads a previous frame's data out of a non-coherent cache."""
    w, h = 640, 480
    frames = [synth.config_frame(w, h, k) for k in range(3)] + [synth.textured_frame(w, h, 77)]
    want = [oracle.detect_describe(f, cap=50000) for f in frames]
    det = clb.Detector(w, h, maxkp=50000)
    for it in range(3):
        for f, (okps, odesc) in zip(frames, want):
            kps, desc = det.detect_and_describe(f)
            assert np.array_equal(kps.view(np.uint8), okps.view(np.uint8)) and np.array_equal(desc, odesc), it
    det.close()
    fs = clb.FrameStream(w, h, depth=2, maxkp=50000)
    order = [0, 1, 2, 3, 3, 1, 0, 2, 2, 0]
    got = []
    for i, k in enumerate(order):
        if fs.in_flight() == 2:
            got.append(fs.next()[:3])
        fs.submit(frames[k], tag=i)
    while fs.in_flight():
        got.append(fs.next()[:3])
    for tag, kps, desc in got:
        okps, odesc = want[order[tag]]
        assert np.array_equal(kps.view(np.uint8), okps.view(np.uint8)) and np.array_equal(desc, odesc), tag
    fs.close()


def test_pinned_outputs_written_by_the_kernel(clb, oracle):
    """page-locked descriptor buffer: the describe kernel writes it directly (no D2H copy); same bytes as the copy path, the
    capacity error still reports the needed count and nothing is written beyond `cap` rows"""
    w, h = 640, 480
    img = synth.config_frame(w, h, 2)
    okps, odesc = oracle.detect_describe(img, cap=50000)
    hold, pimg = clb.pinned_like(img)
    det = clb.Detector(w, h, maxkp=50000)
    for it in range(3):
        kps, desc = det.detect_and_describe(pimg, pinned=True)
        assert np.array_equal(kps.view(np.uint8), okps.view(np.uint8)) and np.array_equal(desc, odesc)
    small = len(okps) - 100
    det._dbuf[small:] = np.uint64(0x5555555555555555)
    with pytest.raises(clb.ClbError) as e:
        det.detect_and_describe(pimg, cap=None, pinned=True) if False else clb.api._check(clb.lib().clb_detect_describe(
            det._h, clb.api._ptr(pimg), w, h, w, clb.api._ptr(det._kbuf), clb.api._ptr(det._dbuf), small, __import__("ctypes").byref(__import__("ctypes").c_int(0))))
    assert e.value.code == 3
    assert (det._dbuf[small:] == np.uint64(0x5555555555555555)).all()
    det.close()


@pytest.mark.parametrize("name", ["frame168.npz", "frame214.npz"])
def test_golden_frame(clb, name):
    """committed fixtures whose keypoints were checked against the reference's compiled KFAST.h + FeatureAngle.h when they were
    generated (tools/gen_golden.py); frame214 has two levels (214 and 86 px wide) on which KFAST.h:243 drops columns"""
    g = np.load(os.path.join(GOLD, name))
    h, w = g["img"].shape
    det = clb.Detector(w, h, maxkp=5000, thresh=int(g["thresh"]))
    kps, desc = det.detect_and_describe(g["img"])
    assert np.array_equal(kps.view(np.uint8), g["kps"].view(np.uint8)) and np.array_equal(desc, g["desc"])
    det.close()


def test_pair_pipeline_recovers_homography(clb, oracle):
    """C1-shaped: detect+describe both views, 2-NN match (margin 40), matches agree with the known homography."""
    w, h = 640, 480
    a, b, H = synth.frame_pair(w, h, 1)
    det = clb.Detector(w, h, maxkp=50000)
    ka, da = det.detect_and_describe(a); ka, da = ka.copy(), da.copy()
    kb, db = det.detect_and_describe(b)
    m = clb.Matcher(maxkp=50000)
    match = m.compute_matches(da, db, 40)
    assert np.array_equal(match, oracle.match_2nn(da, db, 40)[0])
    idx = np.nonzero(match >= 0)[0]
    assert len(idx) > 50
    sa = np.float32(1.2) ** ka["scale"][idx]; sb = np.float32(1.2) ** kb["scale"][match[idx]]
    pa = np.stack([ka["x"][idx] * sa, ka["y"][idx] * sa, np.ones(len(idx))])
    pb = H @ pa; pb = pb[:2] / pb[2]
    err = np.hypot(pb[0] - kb["x"][match[idx]] * sb, pb[1] - kb["y"][match[idx]] * sb)
    assert np.median(err) < 3.0 and (err < 6).mean() > 0.8
    det.close(); m.close()


# ------------------------------------------------------------------ frame stream (throughput mode)
@pytest.mark.parametrize("depth", [1, 3])
def test_frame_stream_equals_single_frame_calls(clb, depth):
    """clb_stream_*: frames of different content and size, pinned and pageable, more frames than ring slots;
    every result bit-identical to clb_detect_describe and returned in submission order."""
    import torch
    sizes = [(640, 480), (333, 251), (640, 480), (512, 384), (640, 480), (64, 48), (640, 480)]
    frames = [synth.textured_frame(w, h, 50 + i) for i, (w, h) in enumerate(sizes)]
    det = clb.Detector(640, 480, maxkp=30000, thresh=30)
    want = []
    for f in frames:
        k, d = det.detect_and_describe(f)
        want.append((k.copy(), d.copy()))
    det.close()
    fs = clb.FrameStream(640, 480, depth=depth, maxkp=30000, thresh=30)
    pinned = []
    got = {}
    nxt = 0
    for i, f in enumerate(frames):
        if i % 2 == 0:                                        # every other frame from pinned memory
            t = torch.empty(f.shape, dtype=torch.uint8).pin_memory(); t.numpy()[:] = f; pinned.append(t); f = t.numpy()
        if fs.in_flight() == depth:
            tag, k, d, dk, dd, n = fs.next()
            assert tag == nxt and dk and dd; got[tag] = (k, d); nxt += 1
        fs.submit(f, tag=i)
    with pytest.raises(clb.ClbError):                          # ring full -> loud error, nothing dropped
        while True:
            fs.submit(frames[0], tag=99)
    while fs.in_flight():
        tag, k, d, dk, dd, n = fs.next()
        if tag == 99:
            continue
        assert tag == nxt; got[tag] = (k, d); nxt += 1
    assert nxt == len(frames)
    for i, (k, d) in enumerate(want):
        assert len(k) > 0 or sizes[i][0] < 100
        assert np.array_equal(got[i][0].view(np.uint8), k.view(np.uint8)) and np.array_equal(got[i][1], d), i
    with pytest.raises(clb.ClbError):
        fs.next()
    fs.close()


@pytest.mark.parametrize("tc", [0, 1])
def test_resident_map_dev_api(clb, oracle, tc):
    """clb_set_map_dev / clb_match_map_dev / clb_match_map_partial_dev: the resident shard is copied (and expanded) once,
    several query sets follow; a new map replaces the cached expansion; partials of two resident shards merge exactly."""
    import torch
    clb.set_param("match_tc", tc)
    try:
        nq = 900
        outs = []
        m = clb.Matcher(maxkp=4096)
        for seed, nt in ((5, 3000), (6, 2177)):
            T = synth.random_descriptors(nt, seed); T[nt - 1] = T[3]
            dT = torch.from_numpy(T.view(np.int64)).cuda()
            m.set_map_dev(dT, nt)
            dT.zero_()                                        # the matcher owns a copy
            for qs in (7, 8):
                Q, _ = synth.planted_queries(T, nq, qs); Q[0] = T[3]
                dQ = torch.from_numpy(Q.view(np.int64)).cuda()
                out = torch.zeros((3, nq), dtype=torch.int32, device="cuda")
                m.match_map_dev(dQ, nq, 60, out[0], out[1], out[2])
                torch.cuda.synchronize()
                ref = oracle.match_2nn(Q, T, 60)
                assert all(np.array_equal(out[i].cpu().numpy(), ref[i]) for i in range(3))
        # two resident shards (two matchers), partial + merge == single scan
        T = synth.random_descriptors(5000, 9); Q, _ = synth.planted_queries(T, nq, 10)
        dQ = torch.from_numpy(Q.view(np.int64)).cuda()
        parts = torch.zeros((2, nq, 4), dtype=torch.int32, device="cuda")
        m2 = clb.Matcher(maxkp=4096)
        for i, (mm, a, b) in enumerate(((m, 0, 2600), (m2, 2600, 5000))):
            mm.set_map_dev(torch.from_numpy(T[a:b].view(np.int64)).cuda(), b - a)
            mm.match_map_partial_dev(dQ, nq, a, parts[i])
        torch.cuda.synchronize()
        out = torch.zeros((3, nq), dtype=torch.int32, device="cuda")
        m.merge_partials_dev(parts, 2, nq, 40, out[0], out[1], out[2])
        torch.cuda.synchronize()
        ref = oracle.match_2nn(Q, T, 40)
        assert all(np.array_equal(out[i].cpu().numpy(), ref[i]) for i in range(3))
        m.close(); m2.close()
    finally:
        clb.set_param("match_tc", -1)


# ------------------------------------------------------------------ device FeatureMap, pair lists, map database
def _ragged_sets(sizes, seed):
    """descriptor sets that share planted rows (so pairs have real matches) plus duplicates (ties)"""
    base = synth.random_descriptors(max(max(sizes), 8), seed)
    sets = []
    for k, n in enumerate(sizes):
        if n == 0:
            sets.append(np.zeros((0, 8), np.uint64)); continue
        S, _ = synth.planted_queries(base, n, seed + 10 + k)
        if n > 6:
            S[n - 1] = S[2]                                   # duplicate rows inside a training set
        sets.append(S)
    return sets


def test_fmap_put_get_erase(clb):
    f = clb.FeatureMap()
    d = synth.random_descriptors(300, 1)
    k = np.zeros(300, clb.KP_DTYPE); k["x"] = np.arange(300); k["angle"] = 0.5
    f.put(7, d, k); f.put(9, d[:5])
    assert len(f) == 2 and 7 in f and 8 not in f and f.count(9) == 5
    kk, dd = f.get(7, want_kps=True)
    assert np.array_equal(dd, d) and np.array_equal(kk, k)
    f.put(7, d[:11])                                          # regions[idx] is replaced, not appended to
    assert f.count(7) == 11 and np.array_equal(f.get(7), d[:11])
    f.put(3, np.zeros((0, 8), np.uint64)); assert f.count(3) == 0 and f.get(3).shape == (0, 8)
    with pytest.raises(clb.ClbError) as e:
        f.count(1234)
    assert e.value.code == 6
    with pytest.raises(clb.ClbError):
        f.get(9, want_kps=True)                               # stored without keypoints
    f.erase(9); assert 9 not in f
    with pytest.raises(clb.ClbError):
        f.erase(9)
    f.close()


@pytest.mark.parametrize("sizes", [(0, 1, 5, 130, 257, 1000), (3000, 2177, 4100), (129, 128, 127, 1)])
def test_match_pairs_vs_oracle(clb, oracle, sizes):
    sets = _ragged_sets(sizes, 40 + len(sizes))
    f = clb.FeatureMap(); m = clb.Matcher(maxkp=16)
    for i, S in enumerate(sets):
        f.put(100 + i, S)
    n = len(sets)
    pairs = [(100 + i, 100 + j) for i in range(n) for j in range(n)]       # includes self pairs (distance 0 / duplicates)
    for thr in (40, 0):
        got = m.match_pairs(f, pairs, thr, want_dist=True)
        for k, (qi, ti) in enumerate(pairs):
            Q, T = sets[qi - 100], sets[ti - 100]
            if len(Q) == 0:
                assert len(got[0][k]) == 0; continue
            if len(T) == 0:
                assert (got[0][k] == -1).all(); continue
            ref = oracle.match_2nn(Q, T, thr)
            for name, a, b in zip(("match", "best", "second"), (got[0][k], got[1][k], got[2][k]), ref):
                assert np.array_equal(a, b), (thr, qi, ti, name, int((a != b).sum()))
    assert clb.get_param("last_match_path") == 4
    m.close(); f.close()


def test_match_pairs_row_slices_and_ratio(clb, oracle):
    """query-row slices (what SURVEY 8e shards over GPUs) concatenate to the full pair; ratio mode equals clb_match_2nn_ratio"""
    sets = _ragged_sets((2500, 1800, 900), 77)
    f = clb.FeatureMap(); m = clb.Matcher(maxkp=4096)
    for i, S in enumerate(sets):
        f.put(i, S)
    full = m.match_pairs(f, [(0, 1), (2, 0)], 40)
    cut = m.match_pairs(f, [(0, 1, 0, 700), (0, 1, 700, 701), (0, 1, 701, 2500), (2, 0, 0, 0), (2, 0, 0, 900)], 40)
    assert np.array_equal(np.concatenate(cut[:3]), full[0]) and len(cut[3]) == 0 and np.array_equal(cut[4], full[1])
    assert np.array_equal(full[0], oracle.match_2nn(sets[0], sets[1], 40)[0])
    r = m.match_pairs(f, [(1, 2)], ratio=0.8, want_dist=True)
    one = m.compute_matches_ratio(sets[1], sets[2], 0.8, want_dist=True)
    assert all(np.array_equal(r[i][0], one[i]) for i in range(3))
    with pytest.raises(clb.ClbError):
        m.match_pairs(f, [(0, 5)], 40)                        # unknown set
    with pytest.raises(clb.ClbError):
        m.match_pairs(f, [(0, 1, 10, 5000)], 40)              # rows outside the set
    m.close(); f.close()


def test_match_pairs_config3_equals_pairwise_calls(clb):
    """config 3 shape: 4 views of one scene, the 12 ordered pairs in one launch == 12 clb_match_2nn calls (themselves oracle-checked above)"""
    import torch
    views, _ = synth.config_views(1280, 720, 4)
    det = clb.Detector(1280, 720, maxkp=120000); m = clb.Matcher(maxkp=30000); f = clb.FeatureMap()
    descs = []
    for k, v in enumerate(views):
        kp, d = det.detect_and_describe(v)
        descs.append(d.copy()); f.put(k, d, kp)
    pairs = [(i, j) for i in range(4) for j in range(4) if i != j]
    got = m.match_pairs(f, pairs, 40)
    for k, (i, j) in enumerate(pairs):
        assert np.array_equal(got[k], m.compute_matches(descs[i], descs[j], 40)), (i, j)
    # device outputs, caller's stream
    tot = m.pair_rows(f, pairs)
    out = torch.full((tot,), -7, dtype=torch.int32, device="cuda")
    m.match_pairs_dev(f, pairs, 40, out, stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), np.concatenate(got))
    det.close(); m.close(); f.close()


def test_fmap_put_dev_from_detector(clb):
    """device hand-off: detect+describe on the device, rows go into the FeatureMap without touching the host"""
    import torch
    img = synth.config_frame(1280, 720, 5)
    det = clb.Detector(1280, 720, maxkp=60000); f = clb.FeatureMap()
    kp, d = det.detect_and_describe(img)
    dk = torch.from_numpy(kp.view(np.uint8).copy()).cuda(); dd = torch.from_numpy(d.view(np.int64).copy()).cuda()
    st = torch.cuda.current_stream().cuda_stream
    f.put_dev(4, dd, len(d), dk, stream=st)
    torch.cuda.synchronize()
    kk, d2 = f.get(4, want_kps=True)
    assert np.array_equal(d2, d) and kk.tobytes() == kp.tobytes()
    det.close(); f.close()


def test_set_map_from_fmap(clb, oracle):
    """setupMapDatabase + setMapData as a device gather: map rows in observation order, UndefinedIndexT skipped"""
    rng = np.random.default_rng(3)
    sets = _ragged_sets((1200, 800, 50), 11)
    f = clb.FeatureMap(); m = clb.Matcher(maxkp=4096)
    for i, S in enumerate(sets):
        f.put(10 * i, S)
    n_obs = 5000
    idx = rng.integers(0, 3, n_obs)
    feat = np.array([rng.integers(0, len(sets[i])) for i in idx], np.int64)
    undef = rng.random(n_obs) < 0.1
    feat[undef] = clb.UNDEFINED_FEAT
    n_map = m.set_map_from_fmap(f, np.stack([idx * 10, feat], 1))
    keep = ~undef
    M = np.stack([sets[i][ft] for i, ft in zip(idx[keep], feat[keep])])
    assert n_map == len(M)
    Q, _ = synth.planted_queries(M, 700, 5)
    got = m.match_with_map(Q, 60, want_dist=True)
    ref = oracle.match_2nn(Q, M, 60)
    assert all(np.array_equal(a, b) for a, b in zip(got, ref))
    assert m.set_map_from_fmap(f, np.zeros((0, 2), np.int64)) == 0
    with pytest.raises(clb.ClbError):
        m.set_map_from_fmap(f, [(10, 800)])                   # feature outside set 10 (800 rows)
    with pytest.raises(clb.ClbError):
        m.set_map_from_fmap(f, [(11, 0)])                     # unknown set
    m.close(); f.close()


@pytest.mark.parametrize("tc", [1, 0])
def test_streamed_host_db_equals_oracle(clb, oracle, tc):
    """DB rows streamed from host memory in slices (clb_match_partial_hostdb, and clb_match_2nn above one slice): the slices fold
    in row order, so duplicates across slice borders still resolve to the first row"""
    import torch
    clb.set_param("match_tc", tc); clb.set_param("match_slice_rows", 1024)
    try:
        nt, nq = 5000, 700
        T = synth.random_descriptors(nt, 31); Q, _ = synth.planted_queries(T, nq, 32)
        T[4100] = T[50]; T[1024] = T[1023]; Q[0] = T[50]; Q[1] = T[1023]     # ties across slice borders
        ref = oracle.match_2nn(Q, T, 40)
        m = clb.Matcher(maxkp=4096)
        got = m.compute_matches(Q, T, 40, want_dist=True)                    # host API, 2 slices (1024 + 3976 rows)
        assert all(np.array_equal(a, b) for a, b in zip(got, ref))
        dQ = torch.from_numpy(Q.view(np.int64)).cuda()
        Tp = torch.from_numpy(T.view(np.int64)).pin_memory()
        part = torch.zeros((nq, 4), dtype=torch.int32, device="cuda"); out = torch.zeros((3, nq), dtype=torch.int32, device="cuda")
        st = torch.cuda.current_stream().cuda_stream
        for _ in range(2):                                                   # second call reuses the slice buffers
            m.match_partial_hostdb(dQ, nq, Tp, nt, 0, part, stream=st)
            m.merge_partials_dev(part, 1, nq, 40, out[0], out[1], out[2], stream=st)
            torch.cuda.synchronize()
            assert all(np.array_equal(out[i].cpu().numpy(), ref[i]) for i in range(3))
        # two host-resident shards with global row bases, merged like the multi-GPU path
        parts = torch.zeros((2, nq, 4), dtype=torch.int32, device="cuda")
        m.match_partial_hostdb(dQ, nq, Tp[:2600], 2600, 0, parts[0], stream=st)
        m.match_partial_hostdb(dQ, nq, Tp[2600:], nt - 2600, 2600, parts[1], stream=st)
        m.merge_partials_dev(parts, 2, nq, 40, out[0], out[1], out[2], stream=st)
        torch.cuda.synchronize()
        assert all(np.array_equal(out[i].cpu().numpy(), ref[i]) for i in range(3))
        m.close()
    finally:
        clb.set_param("match_tc", -1); clb.set_param("match_slice_rows", 131072)


@pytest.mark.parametrize("nq,nt", [(300, 5000), (5000, 5000), (2049, 4100), (777, 12345), (20000, 20000), (130, 70000)])
def test_match_balanced_jobs_vs_oracle(clb, oracle, nq, nt):
    """balanced job layout (query-block strips cut into one range per SM, pieces of unequal length per query block) == oracle,
    for direct calls, partial outputs with a row base, and the resident map"""
    import torch
    clb.set_param("match_tc", 1); clb.set_param("match_balance", 1)
    try:
        T = synth.random_descriptors(nt, 3 + nt); Q, _ = synth.planted_queries(T, nq, 5 + nq)
        T[nt // 2] = T[1]; T[nt - 1] = T[1]; Q[0] = T[1]
        ref = oracle.match_2nn(Q, T, 40)
        m = clb.Matcher(maxkp=4096)
        got = m.compute_matches(Q, T, 40, want_dist=True)
        assert clb.get_param("last_match_balanced") == 1
        assert all(np.array_equal(a, b) for a, b in zip(got, ref))
        dQ = torch.from_numpy(Q.view(np.int64)).cuda(); dT = torch.from_numpy(T.view(np.int64)).cuda()
        cut = nt // 3
        parts = torch.zeros((2, nq, 4), dtype=torch.int32, device="cuda"); out = torch.zeros((3, nq), dtype=torch.int32, device="cuda")
        m.match_partial_dev(dQ, nq, dT[:cut], cut, 0, parts[0]); m.match_partial_dev(dQ, nq, dT[cut:], nt - cut, cut, parts[1])
        m.merge_partials_dev(parts, 2, nq, 40, out[0], out[1], out[2])
        torch.cuda.synchronize()
        assert all(np.array_equal(out[i].cpu().numpy(), ref[i]) for i in range(3))
        m.set_map_data(T)
        got = m.match_with_map(Q, 40, want_dist=True)
        assert all(np.array_equal(a, b) for a, b in zip(got, ref))
        clb.set_param("match_balance", 0)
        got = m.compute_matches(Q, T, 40, want_dist=True)
        assert clb.get_param("last_match_balanced") == 0 and all(np.array_equal(a, b) for a, b in zip(got, ref))
        m.close()
    finally:
        clb.set_param("match_tc", -1); clb.set_param("match_balance", -1)


@pytest.mark.parametrize("flt", [1, 2, 3])
def test_match_filtered_equals_reference(clb, oracle, flt):
    """clb_match_filtered (symmetric cross-check / distance ratio / both over 2-NN lists in both directions, one GPU
    sweep) == matching::Filter: the restatement, and the reference's compiled code when oracle/_ref is there"""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from _helpers import filter_problem
    A, B = filter_problem()
    m = clb.Matcher(maxkp=1024)
    got = m.match_filtered(A, B, flt, 0.8)
    assert np.array_equal(got, oracle.filter_matches(A, B, flt, 0.8))
    if oracle.ref_cpu() is not None:
        assert np.array_equal(got, oracle.ref_filter_matches(A, B, flt, 0.8))
    got = m.match_filtered(B, A, flt, 0.8)                                # the other way round, and a second call on the same matcher
    assert np.array_equal(got, oracle.filter_matches(B, A, flt, 0.8))
    with pytest.raises(clb.ClbError):
        m.match_filtered(A[:1], B, 2, 0.8)
    m.close()

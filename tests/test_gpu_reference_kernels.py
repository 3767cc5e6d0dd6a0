"""On the GPU box the reference's OWN three CUDA kernels (src/CUDALERP.cu, CLATCH.cu, CUDAK2NN.cu, compiled
unmodified into oracle/_ref/libcoloc_ref_gpu.so) run as a second oracle: they pin the C restatement
(oracle/coloc_oracle.c) and, transitively, the new CUDA path (tests/test_gpu_parity.py)."""
import numpy as np
import pytest

from coloc_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def refgpu(oracle, clb):
    if oracle.ref_gpu() is None:
        pytest.skip("oracle/_ref/libcoloc_ref_gpu.so not built")
    return oracle


@pytest.mark.parametrize("nq,nt", [(256, 1000), (1000, 3000), (5000, 5000)])
def test_reference_k2nn_equals_oracle_and_ours(refgpu, clb, nq, nt):
    O = refgpu
    T = synth.random_descriptors(nt, 1); Q, _ = synth.planted_queries(T, nq, 2)
    T[nt // 2] = T[5]; Q[0] = T[5]                          # tie: the first index must win in all three
    for thr in (40, 60):
        ref, _ = O.ref_gpu_k2nn(Q, T, thr)
        assert np.array_equal(ref, O.match_2nn(Q, T, thr)[0])
        m = clb.Matcher(maxkp=nt)
        assert np.array_equal(ref, m.compute_matches(Q, T, thr))
        m.close()


def test_reference_clatch_equals_oracle_and_ours(refgpu, clb):
    O = refgpu
    img = synth.textured_frame(640, 480, 4)
    kps, desc, pyr = O.detect_describe(img, thresh=30, cap=50000, want_pyramid=True)
    assert len(kps) > 2000
    ref, _ = O.ref_gpu_latch(pyr, kps)
    bad = (ref != desc).any(axis=1)
    assert not bad.any(), (int(bad.sum()), len(kps))       # restatement == real CLATCH kernel, bit for bit
    det = clb.Detector(640, 480, maxkp=len(kps) + 1)
    assert np.array_equal(det.describe(pyr, kps), ref)
    det.close()


@pytest.mark.parametrize("w,h", [(640, 480), (1280, 720)])
def test_reference_lerp_exact(refgpu, clb, w, h):
    """north_star: pyramid pixels within 1 LSB of the reference interpolation (texture hardware: 8-bit
    fractional tex2Dgather footprint + unorm8 conversion).  Reports the exact-match fraction too."""
    O = refgpu
    img = synth.textured_frame(w, h, 6)
    ws, hs, fs = O.level_dims(w, h)
    det = clb.Detector(w, h, maxkp=1000)
    ours = det.pyramid(img)
    for L in range(1, 8):
        ref = O.ref_gpu_resize_level(img, fs[L], ws[L], hs[L])
        d = np.abs(ref.astype(int) - ours[L].astype(int))
        assert d.max() <= 1, (L, int(d.max()), float((d > 0).mean()))      # north_star tolerance
        assert d.max() == 0, (L, int(d.max()))                              # observed: bit-exact (8-bit gather footprint rule)
        assert np.array_equal(ref, O.resize_level(img, fs[L], ws[L], hs[L]))  # and so is the C restatement
    det.close()

"""Two-view model residuals (csrc/geometry.cu): the Errors() pass of the AC-RANSAC kernel adaptors for a batch of
hypotheses.  Floating point (fp64): the CUDA kernel, the plain-C restatement (oracle) and the reference's own Eigen functors
(oracle/_ref, compiled from openMVG) agree to RTOL = 1e-9 relative + ATOL = 1e-10 px^2 (they differ only in FMA contraction)."""
import os

import numpy as np
import pytest

import oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
RTOL, ATOL = 1e-9, 1e-10
needs_ref = pytest.mark.skipif(O.ref_cpu() is None or not hasattr(O.ref_cpu(), "ref_model_residuals"), reason="oracle/_ref not built")


def scene(n, seed, outliers=0.3):
    """matched positions of two 1280x720 views related by a homography (+ noise, + gross outliers) and hypothesis batches"""
    rng = np.random.default_rng(seed)
    x1 = rng.uniform([0, 0], [1280, 720], (n, 2))
    H = np.eye(3) + rng.normal(0, [[1e-2, 1e-2, 5], [1e-2, 1e-2, 5], [1e-5, 1e-5, 0]])
    p = np.c_[x1, np.ones(n)] @ H.T
    x2 = p[:, :2] / p[:, 2:] + rng.normal(0, 0.7, (n, 2))
    bad = rng.random(n) < outliers
    x2[bad] = rng.uniform([0, 0], [1280, 720], (int(bad.sum()), 2))
    return x1, x2, H, rng


def hypotheses(kind, H, rng, count):
    if kind == 3:
        return np.stack([H + rng.normal(0, 1e-3, (3, 3)) * np.array([[1, 1, 100], [1, 1, 100], [1e-3, 1e-3, 0]]) for _ in range(count)])
    out = []
    for _ in range(count):
        t = rng.normal(0, 1, 3)
        K = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]])
        out.append(K @ (np.eye(3) + rng.normal(0, 0.05, (3, 3))) * 1e-3)
    return np.stack(out)


@pytest.mark.parametrize("kind", [0, 1, 2, 3])
def test_oracle_matches_golden(kind):
    g = np.load(os.path.join(GOLD, "residuals.npz"))
    got = np.stack([O.model_residuals(kind, M, g["x1"], g["x2"]) for M in g[f"models{kind}"]])
    assert np.allclose(got, g[f"res{kind}"], rtol=RTOL, atol=ATOL)


@needs_ref
@pytest.mark.parametrize("kind", [0, 1, 2, 3])
def test_oracle_matches_reference_functors(kind):
    x1, x2, H, rng = scene(3000, 10 + kind)
    for M in hypotheses(kind, H, rng, 8):
        a = O.model_residuals(kind, M, x1, x2); b = O.ref_model_residuals(kind, M, x1, x2)
        assert np.allclose(a, b, rtol=RTOL, atol=ATOL) and np.isfinite(b).all()


@pytest.mark.gpu
@pytest.mark.parametrize("kind", [0, 1, 2, 3])
@pytest.mark.parametrize("n,nh", [(1, 1), (255, 3), (4097, 64), (20000, 300)])
def test_cuda_residuals_vs_oracle(clb, kind, n, nh):
    x1, x2, H, rng = scene(n, 20 + kind + n)
    Ms = hypotheses(kind, H, rng, nh)
    m = clb.Matcher(maxkp=16)
    thr = 4.0
    res, inl = m.model_residuals(kind, Ms, x1, x2, thr)
    ref = np.stack([O.model_residuals(kind, M, x1, x2) for M in Ms])
    assert np.allclose(res, ref, rtol=RTOL, atol=ATOL)
    lo = (ref < thr * (1 - 1e-8) - ATOL).sum(1); hi = (ref < thr * (1 + 1e-8) + ATOL).sum(1)
    assert ((inl >= lo) & (inl <= hi)).all()
    _, inl2 = m.model_residuals(kind, Ms, x1, x2, thr, want_residuals=False)      # counts only: no residual matrix leaves the GPU
    assert np.array_equal(inl, inl2)
    if O.ref_cpu() is not None and hasattr(O.ref_cpu(), "ref_model_residuals"):
        assert np.allclose(res[0], O.ref_model_residuals(kind, Ms[0], x1, x2), rtol=RTOL, atol=ATOL)
    m.close()


@pytest.mark.gpu
def test_cuda_residuals_edge_cases(clb):
    m = clb.Matcher(maxkp=16)
    x = np.zeros((0, 2)); res, inl = m.model_residuals(0, np.zeros((2, 3, 3)), x, x, 1.0)
    assert res.shape == (2, 0) and (inl == 0).all()
    res, inl = m.model_residuals(3, np.zeros((0, 3, 3)), np.ones((5, 2)), np.ones((5, 2)), 1.0)
    assert res.shape == (0, 5) and inl.shape == (0,)
    with pytest.raises(clb.ClbError):
        m.model_residuals(7, np.eye(3)[None], np.ones((5, 2)), np.ones((5, 2)))
    # the homography that maps x1 onto x2 exactly has zero residuals and every match is an inlier
    x1 = np.random.default_rng(1).uniform(0, 1000, (500, 2)); Hm = np.array([[1.0, 0, 3], [0, 1, -2], [0, 0, 1]])
    res, inl = m.model_residuals(3, Hm[None], x1, x1 + [3, -2], 1e-9)
    assert np.abs(res).max() < 1e-18 and inl[0] == 500
    m.close()

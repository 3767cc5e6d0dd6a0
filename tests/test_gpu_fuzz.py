"""Seeded random shapes against the oracle, bit for bit (the parametrised sizes of test_gpu_parity.py are hand-picked; these are
not): frame geometries with odd widths and heights -- whatever tile, TMA-box and level-width remainders they produce, quirk widths
((w - 38) % 16 == 0, KFAST.h:243) included where they fall -- through the whole detect + describe call, and (nQ, nT) pairs across
the POPC / tensor-core switch and the chunk / tile / query-block borders through the 2-NN entry points."""
import os

import numpy as np
import pytest

from coloc_b200 import synth

pytestmark = pytest.mark.gpu
N_GEO = int(os.environ.get("CLB_FUZZ_GEOMETRIES", "10"))      # more cases on demand: CLB_FUZZ_GEOMETRIES=80 CLB_FUZZ_SHAPES=120
N_SHAPES = int(os.environ.get("CLB_FUZZ_SHAPES", "16"))


def _geometries(n, seed):
    rng = np.random.default_rng(seed)
    out = []
    while len(out) < n:
        w, h = int(rng.integers(40, 1000)), int(rng.integers(40, 760))
        out.append((w, h))
    return out


def _level_widths(w, levels=8, f=1.2):
    ws, s = [], np.float32(1.0)
    for _ in range(levels):
        ws.append(int(np.float32(w) / s)); s = np.float32(s * np.float32(f))
    return ws


@pytest.mark.parametrize("w,h", _geometries(N_GEO, 20261) + [(646, 500), (1286, 90), (257, 41)])
def test_detect_describe_random_geometry(clb, oracle, w, h):
    img = synth.textured_frame(w, h, 1000 + w + h)
    det = clb.Detector(w, h, maxkp=200000)
    for quirk in (1, 0):
        det.set_param("fast_quirk", quirk)
        kps, desc = det.detect_and_describe(img)
        okps, odesc = oracle.detect_describe(img, quirk=quirk, cap=200000)
        assert len(kps) == len(okps), (w, h, quirk, len(kps), len(okps), _level_widths(w))
        assert np.array_equal(kps.view(np.uint8), okps.view(np.uint8)), (w, h, quirk)
        assert np.array_equal(desc, odesc), (w, h, quirk)
    det.close()


def _shapes(n, seed):
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n):
        kind = rng.integers(0, 4)
        if kind == 0:      # small
            out.append((int(rng.integers(1, 600)), int(rng.integers(1, 600))))
        elif kind == 1:    # around the 2^24-comparison switch
            nq = int(rng.integers(200, 6000)); out.append((nq, max(1, (1 << 24) // nq + int(rng.integers(-300, 300)))))
        elif kind == 2:    # tensor-core path, ragged tiles
            out.append((int(rng.integers(2000, 9000)), int(rng.integers(3000, 60000))))
        else:              # few queries, long DB (many chunks)
            out.append((int(rng.integers(1, 300)), int(rng.integers(20000, 150000))))
    return out


@pytest.mark.parametrize("nq,nt", _shapes(N_SHAPES, 7))
def test_match_random_shapes(clb, oracle, nq, nt):
    import torch
    T = synth.random_descriptors(nt, nq + nt)
    Q, _ = synth.planted_queries(T, nq, nq + nt + 1)
    if nt > 3:                      # duplicated rows on purpose: the first must win, second == best on a tie
        T[nt - 1] = T[1]; T[nt // 2] = T[1]; Q[0] = T[1]
    thr = 40 + (nq + nt) % 30
    ref = oracle.match_2nn(Q, T, thr)
    m = clb.Matcher(maxkp=max(nq, nt))
    got = m.compute_matches(Q, T, thr, want_dist=True)
    for name, a, b in zip(("match", "best", "second"), got, ref):
        assert np.array_equal(a, b), (name, nq, nt, int((a != b).sum()))
    dQ = torch.from_numpy(Q.view(np.int64)).cuda(); dT = torch.from_numpy(T.view(np.int64)).cuda()
    out = torch.full((3, nq), -9, dtype=torch.int32, device="cuda")
    m.match_dev(dQ, nq, dT, nt, thr, out[0], out[1], out[2], stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    for i, name in enumerate(("match", "best", "second")):
        assert np.array_equal(out[i].cpu().numpy(), ref[i]), (name, nq, nt)
    m.close()

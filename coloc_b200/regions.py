"""Host side of the path around the GPU kernels: region packing, openMVG .feat / .desc files and the CPU build's
DistanceRatioMatch semantics -- thin ctypes calls into libcoloc_b200.so (csrc/host_regions.cu).

  pack_regions      GPUDetector::detectFeaturesFile's Keypoint -> SIOPointFeature loop (GPUDetector.hpp:167-182)
  save / load       Binary_Regions::Save / Load (binary_regions.hpp:51-66)
  dedup_matches     tail of matching::DistanceRatioMatch (regions_matcher.hpp:166-181)
  distance_ratio_match   what CPUMatcher::computeMatchesPair runs (CPUMatcher.hpp:67-76), 2-NN on the GPU
"""
import ctypes as C
import numpy as np

from .api import lib, _check, _ptr, KP_DTYPE


def pack_regions(kps):
    """(n, 4) float32: x * 1.2^scale, y * 1.2^scale, 7 * 1.2^scale, angle"""
    k = np.ascontiguousarray(kps, KP_DTYPE)
    out = np.zeros((len(k), 4), np.float32)
    _check(lib().clb_pack_regions(_ptr(k), len(k), _ptr(out)))
    return out


def save(feat_path, desc_path, feats, desc):
    d = np.ascontiguousarray(desc, np.uint64).reshape(-1, 8)
    f = None if feats is None else np.ascontiguousarray(feats, np.float32).reshape(-1, 4)
    if f is not None and len(f) != len(d):
        raise ValueError("features and descriptors differ in length")
    _check(lib().clb_regions_save(None if feat_path is None or f is None else str(feat_path).encode(),
                                  None if desc_path is None else str(desc_path).encode(), _ptr(f), _ptr(d), len(d)))


def count(desc_path):
    n = C.c_longlong(0)
    _check(lib().clb_regions_count(str(desc_path).encode(), C.byref(n)))
    return int(n.value)


def load(feat_path, desc_path):
    """-> (feats (n,4) float32 or None, desc (n,8) uint64 or None)"""
    if desc_path is not None:
        cap = count(desc_path)
    else:
        with open(feat_path) as fh:
            cap = sum(1 for _ in fh) + 1
    f = np.zeros((cap, 4), np.float32) if feat_path is not None else None
    d = np.zeros((cap, 8), np.uint64) if desc_path is not None else None
    n = C.c_int(0)
    _check(lib().clb_regions_load(None if feat_path is None else str(feat_path).encode(),
                                  None if desc_path is None else str(desc_path).encode(), _ptr(f), _ptr(d), cap, C.byref(n)))
    n = int(n.value)
    return (None if f is None else f[:n]), (None if d is None else d[:n])


def dedup_matches(match, feats_i, feats_j):
    """match[q] = database row or -1 -> (k, 2) uint32 (i_ = database idx, j_ = query idx) after both dedup passes"""
    m = np.ascontiguousarray(match, np.int32)
    fi = np.ascontiguousarray(feats_i, np.float32); fj = np.ascontiguousarray(feats_j, np.float32)
    out = np.zeros((max(1, int((m >= 0).sum())), 2), np.uint32)
    n = C.c_int(0)
    _check(lib().clb_dedup_matches(_ptr(m), len(m), _ptr(fi), fi.shape[1], _ptr(fj), fj.shape[1], _ptr(out), len(out), C.byref(n)))
    return out[:int(n.value)]


def distance_ratio_match(matcher, desc_i, feats_i, desc_j, feats_j, ratio=0.8):
    """matching::DistanceRatioMatch(ratio, BRUTE_FORCE_HAMMING, regions_I, regions_J): I = database, J = queries"""
    di = np.ascontiguousarray(desc_i, np.uint64).reshape(-1, 8); dj = np.ascontiguousarray(desc_j, np.uint64).reshape(-1, 8)
    fi = np.ascontiguousarray(feats_i, np.float32).reshape(-1, 4); fj = np.ascontiguousarray(feats_j, np.float32).reshape(-1, 4)
    out = np.zeros((max(1, len(dj)), 2), np.uint32)
    n = C.c_int(0)
    _check(lib().clb_distance_ratio_match(matcher._h, _ptr(di), _ptr(fi), len(di), _ptr(dj), _ptr(fj), len(dj), C.c_float(ratio),
                                          _ptr(out), len(out), C.byref(n)))
    return out[:int(n.value)]

"""ctypes bindings for oracle/liboracle.so (C restatement) and oracle/_ref/*.so (the reference's
own code compiled from /root/reference by oracle/Makefile).  TEST INFRASTRUCTURE ONLY."""
import ctypes as C
import os
import subprocess
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
KP_DTYPE = np.dtype([("x", "<i4"), ("y", "<i4"), ("score", "u1"), ("_p0", "u1", 3),
                     ("angle", "<f4"), ("scale", "u1"), ("_p1", "u1", 3)])
assert KP_DTYPE.itemsize == 20

_u8p, _i32p, _u64p, _f32p = (C.POINTER(C.c_uint8), C.POINTER(C.c_int32), C.POINTER(C.c_uint64), C.POINTER(C.c_float))


def build(ref=True):
    """make the oracle (and oracle/_ref when /root/reference is present)."""
    subprocess.check_call(["make", "-s", "-C", HERE, "oracle"])
    if ref and os.path.isdir("/root/reference/include/coloc"):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


def _p(a, t):
    return a.ctypes.data_as(t)


_lib = None
_ref_cpu = None
_ref_gpu = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(HERE, "liboracle.so")
        if not os.path.exists(path):
            build(ref=False)
        _lib = C.CDLL(path)
        _lib.orc_fast_atan2.restype = C.c_float
        _lib.orc_fast_atan2.argtypes = [C.c_float, C.c_float]
        _lib.orc_feature_angle.restype = C.c_float
    return _lib


def ref_cpu():
    """the reference's CPU code (KFAST.h, FeatureAngle.h, openMVG brute force); None if not built."""
    global _ref_cpu
    if _ref_cpu is None:
        path = os.path.join(HERE, "_ref", "libcoloc_ref_cpu.so")
        if not os.path.exists(path):
            return None
        _ref_cpu = C.CDLL(path)
        _ref_cpu.ref_feature_angle.restype = C.c_float
        _ref_cpu.ref_fast_atan2.restype = C.c_float
        _ref_cpu.ref_fast_atan2.argtypes = [C.c_float, C.c_float]
    return _ref_cpu


def ref_gpu():
    """the reference's three CUDA kernels compiled for sm_100a (needs a GPU to call); None if not built."""
    global _ref_gpu
    if _ref_gpu is None:
        path = os.path.join(HERE, "_ref", "libcoloc_ref_gpu.so")
        if not os.path.exists(path):
            return None
        _ref_gpu = C.CDLL(path)
    return _ref_gpu


def set_threads(n):
    lib().orc_set_threads(int(n))


def get_threads():
    return int(lib().orc_get_threads())


# ------------------------------------------------------------------ restatement wrappers
def level_dims(w, h, scale_factor=1.2, levels=8):
    ws = np.zeros(levels, np.int32); hs = np.zeros(levels, np.int32); fs = np.zeros(levels, np.float32)
    lib().orc_level_dims(w, h, C.c_float(scale_factor), levels, _p(ws, _i32p), _p(hs, _i32p), _p(fs, _f32p))
    return ws.tolist(), hs.tolist(), fs


def resize_level(img, f, dw, dh):
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    out = np.zeros((dh, dw), np.uint8)
    lib().orc_resize_level(_p(img, _u8p), w, h, w, C.c_float(float(f)), _p(out, _u8p), dw, dh, dw)
    return out


def pyramid(img, scale_factor=1.2, levels=8):
    h, w = img.shape
    ws, hs, fs = level_dims(w, h, scale_factor, levels)
    out = [np.ascontiguousarray(img, np.uint8)]
    for i in range(1, levels):
        out.append(resize_level(img, fs[i], ws[i], hs[i]))
    return out


def fast_scores(img, thresh=40, quirk=1):
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    sc = np.zeros((h, w), np.uint8)
    lib().orc_fast_scores(_p(img, _u8p), w, h, w, thresh, quirk, _p(sc, _u8p))
    return sc


def fast_nms(img, thresh=40, quirk=1, cap=None):
    """quirk=1 (default) = what the compiled reference does at the KFAST.h:243 `1 << 32` edge; 0 = every column scanned"""
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    cap = cap or (w * h // 4 + 16)
    kps = np.zeros(cap, KP_DTYPE)
    n = lib().orc_fast_nms(_p(img, _u8p), w, h, w, thresh, quirk, kps.ctypes.data_as(C.c_void_p), cap)
    return kps[:min(n, cap)].copy()


def feature_angle(img, x, y):
    img = np.ascontiguousarray(img, np.uint8)
    return float(lib().orc_feature_angle(_p(img, _u8p), int(x), int(y), img.shape[1]))


def fast_atan2(y, x):
    return float(lib().orc_fast_atan2(float(y), float(x)))


def sincos(a):
    s = C.c_float(); c = C.c_float()
    lib().orc_sincos(C.c_float(float(a)), C.byref(s), C.byref(c))
    return s.value, c.value


def latch_describe(levels, kps):
    levels = [np.ascontiguousarray(l, np.uint8) for l in levels]
    n = len(kps)
    L = len(levels)
    ptrs = (C.c_void_p * L)(*[l.ctypes.data for l in levels])
    ws = np.array([l.shape[1] for l in levels], np.int32)
    hs = np.array([l.shape[0] for l in levels], np.int32)
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    desc = np.zeros((n, 8), np.uint64)
    lib().orc_latch_describe(ptrs, _p(ws, _i32p), _p(hs, _i32p), _p(ws, _i32p),
                             kps.ctypes.data_as(C.c_void_p), n, _p(desc, _u64p))
    return desc


def latch_roi(img, kp):
    img = np.ascontiguousarray(img, np.uint8)
    kp = np.ascontiguousarray(kp, KP_DTYPE).reshape(1)
    roi = np.zeros((64, 64), np.uint8)
    lib().orc_latch_roi(_p(img, _u8p), img.shape[1], img.shape[0], img.shape[1], kp.ctypes.data_as(C.c_void_p), _p(roi, _u8p))
    return roi


def hamming512(a, b):
    a = np.ascontiguousarray(a, np.uint64); b = np.ascontiguousarray(b, np.uint64)
    return int(lib().orc_hamming512(_p(a, _u64p), _p(b, _u64p)))


def match_2nn(Q, T, threshold):
    Q = np.ascontiguousarray(Q, np.uint64).reshape(-1, 8); T = np.ascontiguousarray(T, np.uint64).reshape(-1, 8)
    nQ, nT = len(Q), len(T)
    m = np.zeros(nQ, np.int32); b = np.zeros(nQ, np.int32); s = np.zeros(nQ, np.int32)
    lib().orc_match_2nn(_p(Q, _u64p), nQ, _p(T, _u64p), nT, int(threshold), _p(m, _i32p), _p(b, _i32p), _p(s, _i32p))
    return m, b, s


def detect_describe(img, scale_factor=1.2, levels=8, thresh=40, quirk=1, cap=None, want_pyramid=False):
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    cap = cap or (w * h // 8 + 16)
    kps = np.zeros(cap, KP_DTYPE); desc = np.zeros((cap, 8), np.uint64)
    ws, hs, _ = level_dims(w, h, scale_factor, levels)
    pyr = np.zeros(sum(a * b for a, b in zip(ws, hs)), np.uint8) if want_pyramid else None
    n = lib().orc_detect_describe(_p(img, _u8p), w, h, w, C.c_float(scale_factor), levels, thresh, quirk,
                                  kps.ctypes.data_as(C.c_void_p), _p(desc, _u64p), cap,
                                  _p(pyr, _u8p) if want_pyramid else None)
    if want_pyramid:
        lv, off = [], 0
        for a, b in zip(ws, hs):
            lv.append(pyr[off:off + a * b].reshape(b, a)); off += a * b
        return kps[:n].copy(), desc[:n].copy(), lv
    return kps[:n].copy(), desc[:n].copy()


# ------------------------------------------------------------------ reference (CPU) wrappers
def ref_kfast(img, thresh=40, multithread=True, cap=None):
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    # KFAST reads up to 32 bytes past the last scanned pixel (AVX2 loads): pad the buffer
    buf = np.zeros(h * w + 4096, np.uint8); buf[:h * w] = img.ravel()
    cap = cap or (w * h // 4 + 16)
    kps = np.zeros(cap, KP_DTYPE)
    n = ref_cpu().ref_kfast(_p(buf, _u8p), w, h, w, thresh, int(multithread), kps.ctypes.data_as(C.c_void_p), cap)
    return kps[:min(n, cap)].copy()


def ref_feature_angle(img, x, y):
    img = np.ascontiguousarray(img, np.uint8)
    buf = np.zeros(img.size + 64, np.uint8); buf[:img.size] = img.ravel()
    return float(ref_cpu().ref_feature_angle(_p(buf, _u8p), int(x), int(y), img.shape[1]))


def ref_fast_atan2_many(y, x):
    y = np.ascontiguousarray(y, np.float32); x = np.ascontiguousarray(x, np.float32)
    out = np.zeros(y.shape, np.float32)
    ref_cpu().ref_fast_atan2_many(_p(y, _f32p), _p(x, _f32p), _p(out, _f32p), C.c_long(y.size))
    return out


def ref_hamming(a, b):
    a = np.ascontiguousarray(a).view(np.uint8).ravel(); b = np.ascontiguousarray(b).view(np.uint8).ravel()
    return int(ref_cpu().ref_hamming(_p(a, _u8p), _p(b, _u8p), a.size))


def ref_bruteforce_2nn(T, Q, ratio=0.8):
    """openMVG ArrayMatcherBruteForce + NNdistanceRatio.  Returns (idx[nQ,2], dist[nQ,2], ratio_ok[nQ]) or None."""
    T = np.ascontiguousarray(T, np.uint64).reshape(-1, 8); Q = np.ascontiguousarray(Q, np.uint64).reshape(-1, 8)
    nQ = len(Q)
    idx = np.zeros((nQ, 2), np.int32); dist = np.zeros((nQ, 2), np.int32); ok = np.zeros(nQ, np.uint8)
    rc = ref_cpu().ref_bruteforce_2nn(T.ctypes.data_as(_u8p), len(T), Q.ctypes.data_as(_u8p), nQ, C.c_float(ratio),
                                      _p(idx, _i32p), _p(dist, _i32p), _p(ok, _u8p))
    return None if rc else (idx, dist, ok)


def fast_atan2_many(y, x):
    y = np.ascontiguousarray(y, np.float32); x = np.ascontiguousarray(x, np.float32)
    out = np.zeros(y.shape, np.float32)
    lib().orc_fast_atan2_many(_p(y, _f32p), _p(x, _f32p), _p(out, _f32p), C.c_long(y.size))
    return out


def sincos_many(a):
    a = np.ascontiguousarray(a, np.float32)
    s = np.zeros(a.shape, np.float32); c = np.zeros(a.shape, np.float32)
    lib().orc_sincos_many(_p(a, _f32p), _p(s, _f32p), _p(c, _f32p), C.c_long(a.size))
    return s, c


# ------------------------------------------------------------------ reference (GPU) wrappers
def ref_gpu_resize_level(img, f, dw, dh):
    """the reference's CUDALERP kernel (src/CUDALERP.cu) on the GPU."""
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    out = np.zeros((dh, dw), np.uint8)
    rc = ref_gpu().ref_gpu_lerp(_p(img, _u8p), w, h, C.c_float(float(f)), _p(out, _u8p), dw, dh)
    assert rc == 0
    return out


def ref_gpu_latch(levels, kps, iters=1):
    """the reference's CLATCH kernel (src/CLATCH.cu) on the GPU.  Returns (desc, kernel_ms)."""
    levels = [np.ascontiguousarray(l, np.uint8) for l in levels]
    L = len(levels)
    ptrs = (C.c_void_p * L)(*[l.ctypes.data for l in levels])
    ws = np.array([l.shape[1] for l in levels], np.int32); hs = np.array([l.shape[0] for l in levels], np.int32)
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    desc = np.zeros((len(kps), 8), np.uint64)
    ms = C.c_float(0)
    rc = ref_gpu().ref_gpu_latch(ptrs, _p(ws, _i32p), _p(hs, _i32p), L, kps.ctypes.data_as(C.c_void_p), len(kps),
                                 _p(desc, _u64p), int(iters), C.byref(ms))
    assert rc == 0
    return desc, ms.value


def ref_gpu_k2nn(Q, T, threshold, iters=1):
    """the reference's CUDAK2NN kernel (src/CUDAK2NN.cu) on the GPU.  Returns (match, kernel_ms)."""
    Q = np.ascontiguousarray(Q, np.uint64).reshape(-1, 8); T = np.ascontiguousarray(T, np.uint64).reshape(-1, 8)
    m = np.zeros(len(Q), np.int32)
    ms = C.c_float(0)
    rc = ref_gpu().ref_gpu_k2nn(_p(Q, _u64p), len(Q), _p(T, _u64p), len(T), int(threshold), _p(m, _i32p), int(iters), C.byref(ms))
    assert rc == 0
    return m, ms.value


def ref_regions_save(feat_path, desc_path, feats, desc):
    """the reference's Binary_Regions::Save (compiled from its sources); returns True on success"""
    f = np.ascontiguousarray(feats, np.float32).reshape(-1, 4); d = np.ascontiguousarray(desc, np.uint64).reshape(-1, 8)
    return ref_cpu().ref_regions_save(_p(f, _f32p), d.ctypes.data_as(_u8p), len(d), str(feat_path).encode(), str(desc_path).encode()) == 0


def ref_regions_load(feat_path, desc_path, cap):
    f = np.zeros((cap, 4), np.float32); d = np.zeros((cap, 8), np.uint64)
    nd = C.c_int(0)
    n = ref_cpu().ref_regions_load(str(feat_path).encode(), str(desc_path).encode(), _p(f, _f32p), d.ctypes.data_as(_u8p), cap, C.byref(nd))
    return (None, None) if n < 0 else (f[:n], d[:nd.value])


def ref_distance_ratio_match(desc_i, feats_i, desc_j, feats_j, ratio=0.8):
    """the reference's matching::DistanceRatioMatch (regions_matcher.cpp:19-30) on two AKAZE_Binary_Regions; (k, 2) uint32"""
    di = np.ascontiguousarray(desc_i, np.uint64).reshape(-1, 8); dj = np.ascontiguousarray(desc_j, np.uint64).reshape(-1, 8)
    fi = np.ascontiguousarray(feats_i, np.float32).reshape(-1, 4); fj = np.ascontiguousarray(feats_j, np.float32).reshape(-1, 4)
    out = np.zeros((max(1, len(dj)), 2), np.uint32)
    n = ref_cpu().ref_distance_ratio_match(_p(fi, _f32p), di.ctypes.data_as(_u8p), len(di), _p(fj, _f32p), dj.ctypes.data_as(_u8p), len(dj),
                                           C.c_float(ratio), out.ctypes.data_as(C.c_void_p), len(out))
    return out[:n]


def ref_filter_matches(A, B, filter=1, ratio=0.8):
    """the reference's matching::Filter over its own brute-force 2-NN in both directions (compiled openMVG); (k, 2) uint32"""
    A = np.ascontiguousarray(A, np.uint64).reshape(-1, 8); B = np.ascontiguousarray(B, np.uint64).reshape(-1, 8)
    out = np.zeros((max(1, len(A)), 2), np.uint32)
    n = ref_cpu().ref_filter_matches(A.ctypes.data_as(_u8p), len(A), B.ctypes.data_as(_u8p), len(B), int(filter), C.c_float(ratio),
                                     out.ctypes.data_as(C.c_void_p), len(out))
    return None if n < 0 else out[:n]


def filter_matches(A, B, filter=1, ratio=0.8):
    """restatement of matching::Filter on the oracle's 2-NN (first row wins ties), incl. the dropped last index"""
    m01, b01, s01 = match_2nn(A, B, -1)
    m10, _, _ = match_2nn(B, A, -1)
    idx = []
    for i in range(len(m01)):
        sym = m01[i] >= 0 and m10[m01[i]] == i
        rat = np.float32(b01[i]) < np.float32(ratio) * np.float32(s01[i])
        if ((filter & 1) == 0 or sym) and ((filter & 2) == 0 or rat):
            idx.append(i)
    idx = idx[:-1]
    return np.array([(i, m01[i]) for i in idx], np.uint32).reshape(-1, 2)


_f64p = C.POINTER(C.c_double)


def model_residuals(kind, model, x1, x2):
    """restatement of the reference's two-view error functors; model: 3x3 (row-major), x1 / x2: (n, 2) float64"""
    M = np.ascontiguousarray(model, np.float64).reshape(9); a = np.ascontiguousarray(x1, np.float64).reshape(-1, 2)
    b = np.ascontiguousarray(x2, np.float64).reshape(-1, 2)
    out = np.zeros(len(a), np.float64)
    lib().orc_model_residuals(int(kind), _p(M, _f64p), _p(a, _f64p), _p(b, _f64p), len(a), _p(out, _f64p))
    return out


def ref_model_residuals(kind, model, x1, x2):
    """the reference's own functors (compiled from openMVG), same layout"""
    M = np.ascontiguousarray(model, np.float64).reshape(9); a = np.ascontiguousarray(x1, np.float64).reshape(-1, 2)
    b = np.ascontiguousarray(x2, np.float64).reshape(-1, 2)
    out = np.zeros(len(a), np.float64)
    ref_cpu().ref_model_residuals(int(kind), _p(M, _f64p), _p(a, _f64p), _p(b, _f64p), len(a), _p(out, _f64p))
    return out

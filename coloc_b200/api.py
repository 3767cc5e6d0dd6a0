"""ctypes binding of include/coloc_b200.h.  Mirrors the reference's detector / matcher policy
objects (include/coloc/GPUDetector.hpp, GPUMatcher.hpp): same option fields, same method meaning,
errors raised as ClbError(code).  Fails loudly if libcoloc_b200.so is missing -- no fallback."""
import ctypes as C
import os
import subprocess
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
KP_DTYPE = np.dtype([("x", "<i4"), ("y", "<i4"), ("score", "u1"), ("_p0", "u1", 3),
                     ("angle", "<f4"), ("scale", "u1"), ("_p1", "u1", 3)])
PARTIAL_DTYPE = np.dtype([("best_d", "<i4"), ("best_i", "<i4"), ("second_d", "<i4"), ("_pad", "<i4")])
assert KP_DTYPE.itemsize == 20 and PARTIAL_DTYPE.itemsize == 16

OK, EINVAL, ECUDA, ECAPACITY, ENOMEM, EUNSUPPORTED, ENOTFOUND = range(7)
PAIR_DTYPE = np.dtype([("query_idx", "<u4"), ("train_idx", "<u4"), ("q_lo", "<i4"), ("q_hi", "<i4")])
OBS_DTYPE = np.dtype([("idx", "<u4"), ("feat", "<u4")])
UNDEFINED_FEAT = 0xFFFFFFFF


class ClbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"clb error {code}: {msg}")
        self.code = code


class DetectorOpts(C.Structure):
    _fields_ = [("scale_factor", C.c_float), ("scale_levels", C.c_int32), ("width", C.c_int32),
                ("height", C.c_int32), ("maxkp", C.c_int32), ("thresh", C.c_int32)]


class MatcherOpts(C.Structure):
    _fields_ = [("distRatio", C.c_float), ("thresh", C.c_int32), ("maxkp", C.c_int32)]


def lib_path():
    """libcoloc_b200.so (the product); CLB_LIB=lab selects libcoloc_b200_lab.so, the same sources built with -DCLB_LAB
    (superseded kernel generations + variant / timing knobs, for A/B runs and the variant parity tests)."""
    return os.path.join(HERE, "libcoloc_b200_lab.so" if os.environ.get("CLB_LIB") == "lab" else "libcoloc_b200.so")


def build(lab=False):
    """compile the CUDA extension in-tree (nvcc, sm_100a)."""
    subprocess.check_call(["make", "-s", "-j8", "-C", os.path.join(HERE, "csrc")] + (["lab"] if lab else []))


def is_lab():
    return get_param("lab") == 1


_lib = None


def lib():
    global _lib
    if _lib is None:
        p = lib_path()
        if not os.path.exists(p):
            raise ImportError(f"{p} is not built: run `make -C coloc_b200/csrc` (or __graft_entry__.build()); "
                              "coloc_b200 has no CPU fallback")
        L = C.CDLL(p)
        L.clb_strerror.restype = C.c_char_p
        L.clb_last_error.restype = C.c_char_p
        L.clb_launch_count.restype = C.c_int64
        L.clb_last_kernel_ms.restype = C.c_float
        L.clb_last_kernel_ms.argtypes = [C.c_char_p]
        L.clb_set_param.argtypes = [C.c_char_p, C.c_int]
        L.clb_detector_set_param.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        L.clb_stream_set_param.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        L.clb_sharded_set_map.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong]
        L.clb_sharded_set_map_host.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong]
        L.clb_sharded_set_map_shard.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_longlong, C.c_int]
        L.clb_sharded_range.argtypes = [C.c_void_p, C.c_longlong, C.c_int, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        L = lib()
        raise ClbError(rc, f"{L.clb_strerror(rc).decode()} ({L.clb_last_error().decode()})")


def _ptr(a):
    return C.c_void_p(a.ctypes.data) if a is not None else C.c_void_p(0)


def launch_count():
    return int(lib().clb_launch_count())


def set_timing(on):
    lib().clb_set_timing(int(on))


def timeline():
    """[(name, ms)] of the marks recorded since the last call (set_timing(2))"""
    buf = C.create_string_buffer(1 << 16)
    _check(lib().clb_timeline(buf, len(buf)))
    return [(a.rsplit(":", 1)[0], float(a.rsplit(":", 1)[1])) for a in buf.value.decode().split()]


def last_kernel_ms(which):
    return float(lib().clb_last_kernel_ms(which.encode()))


def set_param(name, value):
    _check(lib().clb_set_param(name.encode(), int(value)))


def get_param(name):
    v = C.c_int(0)
    _check(lib().clb_get_param(name.encode(), C.byref(v)))
    return int(v.value)


def microbench_pipes(device=0):
    out = (C.c_double * 8)()
    _check(lib().clb_microbench_pipes(int(device), out))
    k = ["popc_per_clk_sm", "lop3_per_clk_sm", "mixed_popc_per_clk_sm", "mixed_lop3_per_clk_sm", "sm_mhz", "sm_count", "popc_gops", "lop3_gops"]
    return dict(zip(k, [float(v) for v in out]))


def microbench_mma(device=0):
    out = (C.c_double * 6)()
    _check(lib().clb_microbench_mma(int(device), out))
    k = ["cycles_per_mma", "mac_per_clk_sm", "sm_mhz", "tflops", "sm_count", "mmas_per_sm"]
    return dict(zip(k, [float(v) for v in out]))


class PinnedArray:
    """numpy view of page-locked host memory from clb_host_alloc (freed with the object)"""

    def __init__(self, shape, dtype):
        self.dtype = np.dtype(dtype)
        n = int(np.prod(shape)) * self.dtype.itemsize
        self._p = C.c_void_p()
        lib().clb_host_alloc.argtypes = [C.c_size_t, C.c_void_p]
        _check(lib().clb_host_alloc(max(n, 64), C.byref(self._p)))
        buf = (C.c_uint8 * max(n, 1)).from_address(self._p.value)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(np.prod(shape))).reshape(shape)

    def __del__(self):
        try:
            if self._p:
                lib().clb_host_free.argtypes = [C.c_void_p]
                lib().clb_host_free(self._p); self._p = C.c_void_p()
        except Exception:
            pass


def pinned_like(a):
    """a page-locked copy of a numpy array (returns (holder, array): keep the holder alive)"""
    h = PinnedArray(a.shape, a.dtype)
    h.array[...] = a
    return h, h.array


def match_2nn_noalloc(dQ, nQ, dT, nT, threshold, d_match, d_best=None, d_second=None, stream=None):
    """clb_match_2nn_noalloc: allocation-free 2-NN on device pointers (static device scratch; nQ <= 32768).  stream None = legacy default"""
    st = C.c_void_p(0) if stream is None else C.c_void_p(int(stream) if int(stream) != 0 else 1)
    _check(lib().clb_match_2nn_noalloc(_dptr(dQ), int(nQ), _dptr(dT), int(nT), int(threshold), _dptr(d_match), _dptr(d_best), _dptr(d_second), st))


def _sarg(stream):
    """cudaStream_t argument: None -> NULL (= the object's own stream); a torch stream handle of 0 (the legacy
    default stream) is passed as cudaStreamLegacy (0x1) so that it is not mistaken for NULL."""
    if stream is None:
        return C.c_void_p(0)
    return C.c_void_p(int(stream) if int(stream) != 0 else 1)


def _dptr(t):
    """device pointer of a torch CUDA tensor (plumbing only); ints pass through."""
    if t is None:
        return C.c_void_p(0)
    if isinstance(t, int):
        return C.c_void_p(t)
    return C.c_void_p(t.data_ptr())


class Detector:
    """GPUDetector (include/coloc/GPUDetector.hpp:29-292) over the C ABI."""

    def __init__(self, width, height, scale_factor=1.2, scale_levels=8, maxkp=5000, thresh=40, device=0, fast_quirk=1):
        """fast_quirk=1 (default): FAST scans the columns the compiled reference scans (KFAST.h:243 `1 << 32` edge on 214- /
        1286-wide levels); 0: every column (include/coloc_b200.h, clb_detector_set_param)."""
        self.opts = DetectorOpts(scale_factor, scale_levels, width, height, maxkp, thresh)
        self._h = C.c_void_p()
        _check(lib().clb_detector_create(C.byref(self.opts), int(device), C.byref(self._h)))
        self.kps = np.zeros(0, KP_DTYPE)      # last frame's raw output, as GPUDetector.hpp:32-33
        self.desc = np.zeros((0, 8), np.uint64)
        if fast_quirk != 1:
            self.set_param("fast_quirk", fast_quirk)

    def set_param(self, name, value):
        _check(lib().clb_detector_set_param(self._h, name.encode(), int(value)))

    def close(self):
        if self._h:
            lib().clb_detector_destroy(self._h)
            self._h = C.c_void_p()

    freeGPUMemory = close

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def level_dims(self, w, h):
        n = self.opts.scale_levels
        ws = np.zeros(n, np.int32); hs = np.zeros(n, np.int32); fs = np.zeros(n, np.float32)
        _check(lib().clb_detector_level_dims(self._h, w, h, _ptr(ws), _ptr(hs), _ptr(fs)))
        return ws.tolist(), hs.tolist(), fs

    def detect_and_describe(self, img, cap=None, copy=True, pinned=False):
        """GPUDetector::detectAndDescribe (GPUDetector.hpp:216-291) on a host uint8 image.  pinned=True keeps the output
        buffers in page-locked memory: the describe kernel then writes the descriptors straight into them."""
        img = np.ascontiguousarray(img, np.uint8)
        h, w = img.shape
        cap = cap if cap is not None else self.opts.maxkp
        if getattr(self, "_cap", -1) != cap or getattr(self, "_pinned", False) != pinned:      # output buffers are kept across calls
            if pinned:
                self._kh = PinnedArray((cap,), KP_DTYPE); self._dh = PinnedArray((cap, 8), np.uint64)
                self._kbuf, self._dbuf = self._kh.array, self._dh.array
            else:
                self._kbuf = np.empty(cap, KP_DTYPE); self._dbuf = np.empty((cap, 8), np.uint64)
            self._cap = cap; self._pinned = pinned
        n = C.c_int(0)
        _check(lib().clb_detect_describe(self._h, _ptr(img), w, h, img.strides[0], _ptr(self._kbuf), _ptr(self._dbuf), cap, C.byref(n)))
        # like the reference's public `kps` / `desc` members these hold the LAST frame: copy=False hands out views that the next call overwrites
        if copy:
            self.kps, self.desc = self._kbuf[:n.value].copy(), self._dbuf[:n.value].copy()
        else:
            self.kps, self.desc = self._kbuf[:n.value], self._dbuf[:n.value]
        return self.kps, self.desc

    def detect_and_describe_dev(self, d_img, stream=None):
        """device-resident variant: torch uint8 CUDA tensor in, (d_kps ptr, d_desc ptr, n) out."""
        h, w = d_img.shape
        pk = C.c_void_p(); pd = C.c_void_p(); n = C.c_int(0)
        _check(lib().clb_detect_describe_dev(self._h, _dptr(d_img), w, h, d_img.stride(0), C.byref(pk), C.byref(pd), C.byref(n),
                                             _sarg(stream)))
        return pk.value, pd.value, n.value

    def pyramid(self, img):
        img = np.ascontiguousarray(img, np.uint8)
        h, w = img.shape
        ws, hs, _ = self.level_dims(w, h)
        out = np.zeros(sum(a * b for a, b in zip(ws, hs)), np.uint8)
        _check(lib().clb_pyramid(self._h, _ptr(img), w, h, img.strides[0], _ptr(out)))
        lv, off = [], 0
        for a, b in zip(ws, hs):
            lv.append(out[off:off + a * b].reshape(b, a)); off += a * b
        return lv

    def fast_nms_angle(self, level, level_idx=0, cap=None):
        level = np.ascontiguousarray(level, np.uint8)
        h, w = level.shape
        cap = cap if cap is not None else self.opts.maxkp
        kps = np.zeros(cap, KP_DTYPE); n = C.c_int(0)
        _check(lib().clb_fast_nms_angle(self._h, _ptr(level), w, h, level.strides[0], int(level_idx), _ptr(kps), cap, C.byref(n)))
        return kps[:n.value]

    def describe(self, levels, kps):
        levels = [np.ascontiguousarray(l, np.uint8) for l in levels]
        L = len(levels)
        ptrs = (C.c_void_p * L)(*[l.ctypes.data for l in levels])
        ws = np.array([l.shape[1] for l in levels], np.int32); hs = np.array([l.shape[0] for l in levels], np.int32)
        kps = np.ascontiguousarray(kps, KP_DTYPE)
        desc = np.zeros((len(kps), 8), np.uint64)
        _check(lib().clb_describe(self._h, ptrs, _ptr(ws), _ptr(hs), _ptr(ws), L, _ptr(kps), len(kps), _ptr(desc)))
        return desc

    def sincos(self, angles):
        a = np.ascontiguousarray(angles, np.float32)
        s = np.zeros_like(a); c = np.zeros_like(a)
        _check(lib().clb_sincos(self._h, _ptr(a), a.size, _ptr(s), _ptr(c)))
        return s, c


class Matcher:
    """GPUMatcher (include/coloc/GPUMatcher.hpp:45-272) over the C ABI."""

    def __init__(self, distRatio=0.8, thresh=60, maxkp=5000, device=0):
        self.opts = MatcherOpts(distRatio, thresh, maxkp)
        self._h = C.c_void_p()
        _check(lib().clb_matcher_create(C.byref(self.opts), int(device), C.byref(self._h)))

    def close(self):
        if self._h:
            lib().clb_matcher_destroy(self._h)
            self._h = C.c_void_p()

    freeGPUMemory = close

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @staticmethod
    def _rows(a):
        a = np.ascontiguousarray(a, np.uint64).reshape(-1, 8)
        return a, len(a)

    def _outs(self, nQ, want_dist):
        m = np.zeros(nQ, np.int32)
        b = np.zeros(nQ, np.int32) if want_dist else None
        s = np.zeros(nQ, np.int32) if want_dist else None
        return m, b, s

    def compute_matches(self, Q, T, threshold=40, want_dist=False):
        """GPUMatcher::computeMatches(hQ, hT, nQ, nT, threshold) (GPUMatcher.hpp:180-226): int match[nQ]."""
        Q, nQ = self._rows(Q); T, nT = self._rows(T)
        m, b, s = self._outs(nQ, want_dist)
        _check(lib().clb_match_2nn(self._h, _ptr(Q), nQ, _ptr(T), nT, int(threshold), _ptr(m), _ptr(b), _ptr(s)))
        return (m, b, s) if want_dist else m

    def compute_matches_ratio(self, Q, T, ratio=None, want_dist=False):
        """ratio-test mode of the CPU build (CPUMatcher.hpp:67-76, matching_filters.hpp:39-60)."""
        Q, nQ = self._rows(Q); T, nT = self._rows(T)
        m, b, s = self._outs(nQ, want_dist)
        r = self.opts.distRatio if ratio is None else ratio
        _check(lib().clb_match_2nn_ratio(self._h, _ptr(Q), nQ, _ptr(T), nT, C.c_float(r), _ptr(m), _ptr(b), _ptr(s)))
        return (m, b, s) if want_dist else m

    def match_filtered(self, A, B, filter=1, ratio=0.8):
        """matching::Filter (matching_filters.hpp:118-213): 1 symmetric cross-check, 2 distance ratio, 3 both -> (k, 2) uint32"""
        A, nA = self._rows(A); B, nB = self._rows(B)
        out = np.zeros((max(nA, 1), 2), np.uint32); n = C.c_int(0)
        _check(lib().clb_match_filtered(self._h, _ptr(A), nA, _ptr(B), nB, int(filter), C.c_float(ratio), _ptr(out), len(out), C.byref(n)))
        return out[:n.value].copy()

    def set_map_data(self, M):
        """GPUMatcher::setMapData (GPUMatcher.hpp:110-117)."""
        M, nM = self._rows(M)
        _check(lib().clb_set_map(self._h, _ptr(M), nM))

    def match_with_map(self, Q, threshold=-1, want_dist=False):
        """setQueryImage + matchFeaturesWithMap (GPUMatcher.hpp:130-141,252-271); returns MAP row per query."""
        Q, nQ = self._rows(Q)
        m, b, s = self._outs(nQ, want_dist)
        _check(lib().clb_match_map(self._h, _ptr(Q), nQ, int(threshold), _ptr(m), _ptr(b), _ptr(s)))
        return (m, b, s) if want_dist else m

    # ---- device-pointer entry points (torch CUDA tensors as plain buffers) -----------------
    def match_dev(self, dQ, nQ, dT, nT, threshold, d_match, d_best=None, d_second=None, stream=None):
        _check(lib().clb_match_2nn_dev(self._h, _dptr(dQ), int(nQ), _dptr(dT), int(nT), int(threshold),
                                       _dptr(d_match), _dptr(d_best), _dptr(d_second), _sarg(stream)))

    def match_partial_dev(self, dQ, nQ, dT, nT, t_base, d_partial, stream=None):
        _check(lib().clb_match_partial_dev(self._h, _dptr(dQ), int(nQ), _dptr(dT), int(nT), C.c_int64(int(t_base)),
                                           _dptr(d_partial), _sarg(stream)))

    def match_partial_hostdb(self, dQ, nQ, T_host, nT, t_base, d_partial, stream=None):
        """DB rows in (pinned) HOST memory streamed in slices behind the sweep; T_host: numpy array, torch CPU tensor or address"""
        if isinstance(T_host, np.ndarray):
            tp = C.c_void_p(T_host.ctypes.data)
        elif isinstance(T_host, int):
            tp = C.c_void_p(T_host)
        else:
            tp = C.c_void_p(T_host.data_ptr())
        _check(lib().clb_match_partial_hostdb(self._h, _dptr(dQ), int(nQ), tp, int(nT), C.c_int64(int(t_base)), _dptr(d_partial), _sarg(stream)))

    def set_map_dev(self, dM, nM, stream=None):
        """device-resident map / DB shard (copied D2D once, expanded once for the tensor-core path)"""
        _check(lib().clb_set_map_dev(self._h, _dptr(dM), int(nM), _sarg(stream)))

    def match_map_dev(self, dQ, nQ, threshold, d_match, d_best=None, d_second=None, stream=None):
        _check(lib().clb_match_map_dev(self._h, _dptr(dQ), int(nQ), int(threshold), _dptr(d_match), _dptr(d_best), _dptr(d_second),
                                       _sarg(stream)))

    def match_map_partial_dev(self, dQ, nQ, t_base, d_partial, stream=None):
        _check(lib().clb_match_map_partial_dev(self._h, _dptr(dQ), int(nQ), C.c_int64(int(t_base)), _dptr(d_partial), _sarg(stream)))

    def merge_partials_dev(self, d_parts, nparts, nQ, threshold, d_match, d_best=None, d_second=None, stream=None):
        _check(lib().clb_merge_partials_dev(self._h, _dptr(d_parts), int(nparts), int(nQ), int(threshold),
                                            _dptr(d_match), _dptr(d_best), _dptr(d_second), _sarg(stream)))


    # ---- device FeatureMap: pair lists and the map database ---------------------
    @staticmethod
    def _pairs(pairs):
        """[(query_idx, train_idx)] or [(query_idx, train_idx, q_lo, q_hi)] -> clb_pair array"""
        a = np.zeros(len(pairs), PAIR_DTYPE)
        for k, p in enumerate(pairs):
            a[k] = (p[0], p[1], p[2] if len(p) > 2 else 0, p[3] if len(p) > 3 else -1)
        return a

    def pair_rows(self, fmap, pairs):
        a = self._pairs(pairs)
        n = C.c_longlong(0)
        _check(lib().clb_match_pairs_rows(self._h, fmap._h, _ptr(a), len(a), C.byref(n)))
        return int(n.value)

    def match_pairs(self, fmap, pairs, threshold=40, ratio=0.0, want_dist=False):
        """GPUMatcher::computeMatches(FeatureMap&, PairWiseMatches&) (GPUMatcher.hpp:143-155) over device-resident sets, the
        whole list in one launch.  Returns one match[] array per pair (and best / second when want_dist)."""
        a = self._pairs(pairs)
        tot = self.pair_rows(fmap, pairs)
        m, b, s = self._outs(tot, want_dist)
        _check(lib().clb_match_pairs(self._h, fmap._h, _ptr(a), len(a), int(threshold), C.c_float(ratio), _ptr(m), _ptr(b), _ptr(s)))
        cuts, off = [], 0
        for p in a:
            n = (fmap.count(int(p["query_idx"])) if p["q_hi"] < 0 else int(p["q_hi"])) - int(p["q_lo"])
            cuts.append((off, off + n)); off += n
        split = lambda v: [v[x:y] for x, y in cuts]
        return (split(m), split(b), split(s)) if want_dist else split(m)

    def match_pairs_dev(self, fmap, pairs, threshold, d_match, d_best=None, d_second=None, ratio=0.0, stream=None):
        a = pairs if isinstance(pairs, np.ndarray) and pairs.dtype == PAIR_DTYPE else self._pairs(pairs)
        _check(lib().clb_match_pairs_dev(self._h, fmap._h, _ptr(a), len(a), int(threshold), C.c_float(ratio),
                                         _dptr(d_match), _dptr(d_best), _dptr(d_second), _sarg(stream)))

    def model_residuals(self, kind, models, x1, x2, threshold=0.0, want_residuals=True):
        """Errors() of the AC-RANSAC kernel adaptors for a batch of hypotheses (clb_model_residuals).
        models: (H, 3, 3) row-major float64; x1, x2: (n, 2) float64.  Returns (residuals (H, n) or None, inlier counts (H,))."""
        M = np.ascontiguousarray(models, np.float64).reshape(-1, 9)
        a = np.ascontiguousarray(x1, np.float64).reshape(-1, 2); b = np.ascontiguousarray(x2, np.float64).reshape(-1, 2)
        if len(a) != len(b):
            raise ValueError("x1 and x2 differ in length")
        res = np.zeros((len(M), len(a)), np.float64) if want_residuals else None
        inl = np.zeros(len(M), np.int32)
        _check(lib().clb_model_residuals(self._h, int(kind), _ptr(M), len(M), _ptr(a), _ptr(b), len(a), C.c_double(threshold), _ptr(res), _ptr(inl)))
        return res, inl

    def model_residuals_dev(self, kind, d_models, n_models, d_x1, d_x2, n, threshold, d_res=None, d_inl=None, stream=None):
        _check(lib().clb_model_residuals_dev(self._h, int(kind), _dptr(d_models), int(n_models), _dptr(d_x1), _dptr(d_x2), int(n),
                                             C.c_double(threshold), _dptr(d_res), _dptr(d_inl), _sarg(stream)))

    def set_map_from_fmap(self, fmap, obs, stream=None):
        """colocData::setupMapDatabase (colocData.hpp:89-121) + setMapData as a device gather: obs = [(set idx, feature id)],
        entries with feature id UNDEFINED_FEAT are skipped.  Returns the number of map rows."""
        a = np.ascontiguousarray(np.asarray(obs, dtype=np.int64).reshape(-1, 2).astype(np.uint32)).view(OBS_DTYPE).reshape(-1)
        n = C.c_int(0)
        _check(lib().clb_set_map_from_fmap(self._h, fmap._h, _ptr(a), len(a), C.byref(n), _sarg(stream)))
        return int(n.value)


class ShardedMatcher:
    """DB-row-sharded map matching over the C ABI (clb_sharded_*): GPUMatcher::setMapData + matchSceneWithMap /
    matchFeaturesWithMap (GPUMatcher.hpp:110-117, 174-178, 252-271) with the map rows cut over several GPUs.
      ShardedMatcher(devices=[0, 1, ...])                 one process drives the listed devices (peer copies of the partials)
      ShardedMatcher(device=d, rank=r, nranks=n, uid=b)   one process per GPU; uid = ShardedMatcher.unique_id() of rank 0,
                                                          broadcast by the caller; the exchange is one ncclAllGather per call"""

    def __init__(self, devices=None, device=None, rank=0, nranks=1, uid=None, distRatio=0.8, thresh=60, maxkp=5000):
        self.opts = MatcherOpts(distRatio, thresh, maxkp)
        self._h = C.c_void_p()
        if devices is not None:
            devs = (C.c_int * len(devices))(*[int(d) for d in devices])
            _check(lib().clb_sharded_create(C.byref(self.opts), devs, len(devices), C.byref(self._h)))
            self.rank_mode = False
        else:
            buf = C.create_string_buffer(bytes(uid), 128) if uid is not None else None
            _check(lib().clb_sharded_create_rank(C.byref(self.opts), int(device or 0), int(rank), int(nranks), buf, C.byref(self._h)))
            self.rank_mode = True
        self.n = int(lib().clb_sharded_shards(self._h))
        self._keep = None

    @staticmethod
    def unique_id():
        buf = C.create_string_buffer(128)
        _check(lib().clb_sharded_unique_id(buf))
        return bytes(buf.raw)

    def close(self):
        if self._h:
            lib().clb_sharded_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def shard_range(self, n_rows, shard):
        lo, hi = C.c_longlong(0), C.c_longlong(0)
        _check(lib().clb_sharded_range(self._h, int(n_rows), int(shard), C.byref(lo), C.byref(hi)))
        return int(lo.value), int(hi.value)

    @staticmethod
    def _host_ptr(a):
        if isinstance(a, np.ndarray):
            return C.c_void_p(a.ctypes.data)
        if isinstance(a, int):
            return C.c_void_p(a)
        return C.c_void_p(a.data_ptr())

    def set_map(self, M, host_resident=False):
        """the WHOLE map (numpy uint64 [n, 8] or pinned torch tensor); host_resident: rows stay in host memory and stream per call"""
        n = int(M.shape[0])
        if host_resident:
            self._keep = M
            _check(lib().clb_sharded_set_map_host(self._h, self._host_ptr(M), n))
        else:
            _check(lib().clb_sharded_set_map(self._h, self._host_ptr(M), n))

    def set_map_shard(self, rows, row_lo, n_total, where=0):
        """rank mode: this rank's rows only.  where: 0 host -> resident, 1 device -> resident, 2 host rows streamed per call"""
        if where == 2:
            self._keep = rows
        ptr = _dptr(rows) if where == 1 else self._host_ptr(rows)
        _check(lib().clb_sharded_set_map_shard(self._h, ptr, int(row_lo), int(rows.shape[0]), int(n_total), int(where)))

    def match_map(self, Q, threshold=-1, want_dist=False, out=None):
        """HOST queries -> HOST match[] (global map row or -1); blocking; collective in rank mode"""
        if isinstance(Q, np.ndarray):
            Q = np.ascontiguousarray(Q, np.uint64).reshape(-1, 8)
        nQ = int(Q.shape[0])
        m = out if out is not None else np.zeros(nQ, np.int32)
        b = np.zeros(nQ, np.int32) if want_dist else None
        s = np.zeros(nQ, np.int32) if want_dist else None
        _check(lib().clb_sharded_match_map(self._h, self._host_ptr(Q), nQ, int(threshold), self._host_ptr(m), _ptr(b), _ptr(s)))
        return (m, b, s) if want_dist else m

    def match_map_dev(self, dQ, nQ, threshold, d_match, d_best=None, d_second=None, stream=None):
        _check(lib().clb_sharded_match_map_dev(self._h, _dptr(dQ), int(nQ), int(threshold), _dptr(d_match), _dptr(d_best), _dptr(d_second),
                                               _sarg(stream)))


class FeatureMap:
    """device-side coloc::FeatureMap (idx -> Regions; GPUDetector.hpp:167-182): every frame's descriptors (and keypoints)
    stay on the GPU between detect and match (clb_fmap_*)."""

    def __init__(self, device=0):
        self._h = C.c_void_p()
        _check(lib().clb_fmap_create(int(device), C.byref(self._h)))

    def close(self):
        if self._h:
            lib().clb_fmap_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __len__(self):
        return int(lib().clb_fmap_size(self._h))

    def put(self, idx, desc, kps=None):
        d = np.ascontiguousarray(desc, np.uint64).reshape(-1, 8)
        k = None
        if kps is not None:
            k = np.ascontiguousarray(kps, KP_DTYPE)
            if len(k) != len(d):
                raise ValueError("keypoints and descriptors differ in length")
        _check(lib().clb_fmap_put(self._h, C.c_uint32(idx), _ptr(k), _ptr(d), len(d)))

    def put_dev(self, idx, d_desc, n, d_kps=None, stream=None):
        _check(lib().clb_fmap_put_dev(self._h, C.c_uint32(idx), _dptr(d_kps), _dptr(d_desc), int(n), _sarg(stream)))

    def count(self, idx):
        n = C.c_int(0)
        _check(lib().clb_fmap_count(self._h, C.c_uint32(idx), C.byref(n)))
        return int(n.value)

    def __contains__(self, idx):
        n = C.c_int(0)
        return lib().clb_fmap_count(self._h, C.c_uint32(idx), C.byref(n)) == 0

    def erase(self, idx):
        _check(lib().clb_fmap_erase(self._h, C.c_uint32(idx)))

    def get(self, idx, want_kps=False):
        n = self.count(idx)
        d = np.zeros((n, 8), np.uint64)
        k = np.zeros(n, KP_DTYPE) if want_kps else None
        got = C.c_int(0)
        _check(lib().clb_fmap_get(self._h, C.c_uint32(idx), _ptr(k), _ptr(d), n, C.byref(got)))
        return (k, d) if want_kps else d

    def desc_dev(self, idx):
        """(device pointer of the packed rows, device pointer of the keypoints or 0, rows)"""
        pd, pk, n = C.c_void_p(0), C.c_void_p(0), C.c_int(0)
        _check(lib().clb_fmap_desc_dev(self._h, C.c_uint32(idx), C.byref(pd), C.byref(pk), C.byref(n)))
        return int(pd.value or 0), int(pk.value or 0), int(n.value)


class FrameStream:
    """throughput mode of the detector (clb_stream_*): up to `depth` frames in flight, results in submission order.
    Same output as Detector.detect_and_describe, frame for frame."""

    def __init__(self, width, height, depth=3, scale_factor=1.2, scale_levels=8, maxkp=5000, thresh=40, device=0, fast_quirk=1):
        self.opts = DetectorOpts(scale_factor, scale_levels, width, height, maxkp, thresh)
        self.depth = int(depth)
        self._h = C.c_void_p()
        L = lib()
        L.clb_stream_submit.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_uint64]
        _check(L.clb_stream_create(C.byref(self.opts), int(device), self.depth, C.byref(self._h)))
        if fast_quirk != 1:
            _check(L.clb_stream_set_param(self._h, b"fast_quirk", int(fast_quirk)))

    def close(self):
        if self._h:
            lib().clb_stream_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def in_flight(self):
        return int(lib().clb_stream_in_flight(self._h))

    def submit(self, img, tag=0):
        """img: contiguous-row uint8 numpy array (pinned or pageable) or (ptr, w, h, stride)"""
        if isinstance(img, tuple):
            ptr, w, h, stride = img
        else:
            assert img.dtype == np.uint8 and img.strides[1] == 1
            h, w = img.shape
            ptr, stride = img.ctypes.data, img.strides[0]
        _check(lib().clb_stream_submit(self._h, C.c_void_p(ptr), int(w), int(h), int(stride), C.c_uint64(int(tag))))

    def next(self, copy=True, want_host=True):
        """-> (tag, kps, desc, d_kps_ptr, d_desc_ptr); kps/desc are views of pinned memory unless copy=True"""
        tag = C.c_uint64(0); n = C.c_int(0)
        pk = C.c_void_p(); pd = C.c_void_p(); dk = C.c_void_p(); dd = C.c_void_p()
        _check(lib().clb_stream_next(self._h, C.byref(tag), C.byref(pk) if want_host else None, C.byref(pd) if want_host else None,
                                     C.byref(dk), C.byref(dd), C.byref(n)))
        kps = desc = None
        if want_host:
            if n.value > 0:
                kps = np.ctypeslib.as_array(C.cast(pk, C.POINTER(C.c_uint8)), shape=(n.value * 20,)).view(KP_DTYPE)
                desc = np.ctypeslib.as_array(C.cast(pd, C.POINTER(C.c_uint64)), shape=(n.value, 8))
                if copy:
                    kps, desc = kps.copy(), desc.copy()
            else:
                kps, desc = np.zeros(0, KP_DTYPE), np.zeros((0, 8), np.uint64)
        return int(tag.value), kps, desc, dk.value, dd.value, int(n.value)

"""GPU probe: per-call latency of clb_detect_describe with pageable vs pinned host buffers (config-2 frame)."""
import os, sys, time, ctypes as C
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import coloc_b200 as clb
from coloc_b200 import synth
from coloc_b200.api import lib, _check

w, h = 1280, 720
img = synth.config_frame(w, h, 0)
det = clb.Detector(w, h, maxkp=60000)
cap = 60000
L = lib()

def run(img_a, kps_a, desc_a, reps=200):
    n = C.c_int(0)
    for _ in range(5):
        _check(L.clb_detect_describe(det._h, C.c_void_p(img_a.ctypes.data), w, h, w, C.c_void_p(kps_a.ctypes.data), C.c_void_p(desc_a.ctypes.data), cap, C.byref(n)))
    t0 = time.perf_counter()
    for _ in range(reps):
        _check(L.clb_detect_describe(det._h, C.c_void_p(img_a.ctypes.data), w, h, w, C.c_void_p(kps_a.ctypes.data), C.c_void_p(desc_a.ctypes.data), cap, C.byref(n)))
    return (time.perf_counter() - t0) / reps * 1e3, n.value

kps = np.zeros(cap, clb.KP_DTYPE); desc = np.zeros((cap, 8), np.uint64)
ms, n = run(img, kps, desc)
print(f"pageable in/out: {ms:.3f} ms/frame ({1e3 / ms:.0f} fps), {n} keypoints")
pimg = torch.from_numpy(img.copy()).pin_memory().numpy()
pk = torch.zeros(cap * 20, dtype=torch.uint8).pin_memory().numpy().view(clb.KP_DTYPE)
pd = torch.zeros((cap, 8), dtype=torch.int64).pin_memory().numpy().view(np.uint64)
ms, n = run(pimg, pk, pd)
print(f"pinned in/out:   {ms:.3f} ms/frame ({1e3 / ms:.0f} fps)")
ms, n = run(pimg, kps, desc)
print(f"pinned in, pageable out: {ms:.3f} ms/frame")
ms, n = run(img, pk, pd)
print(f"pageable in, pinned out: {ms:.3f} ms/frame")
clb.set_timing(True)
run(pimg, pk, pd, reps=3)
print("kernels:", {k: round(clb.last_kernel_ms(k), 4) for k in ("resize", "fast", "describe")})

"""per-stage kernel times (CUDA events inside the library) of detect+describe on the BASELINE frames: C1 640x480, C2 1280x720,
C5 3840x2160 (config_frame k=0), plus the blocking per-frame call latency.  python tools/stage_probe.py"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import coloc_b200 as clb
from coloc_b200 import synth
for (w, h) in ((640, 480), (1280, 720), (3840, 2160)):
    img = torch.from_numpy(synth.config_frame(w, h, 0)).pin_memory().numpy()
    det = clb.Detector(w, h, maxkp=400000)
    clb.set_timing(True)
    ts = []
    for _ in range(12):
        k, d = det.detect_and_describe(img, copy=False)
        ts.append([clb.last_kernel_ms(x) for x in ("resize", "fast", "describe")])
    t = np.min(np.array(ts[2:]), axis=0) * 1e3
    clb.set_timing(False)
    lat = []
    det.detect_and_describe(img, copy=False, pinned=True)
    for _ in range(30):
        t0 = time.perf_counter(); det.detect_and_describe(img, copy=False, pinned=True); lat.append(time.perf_counter() - t0)
    print(f"{w}x{h}: {len(k)} kp  resize {t[0]:.1f} us  fast {t[1]:.1f} us  describe {t[2]:.1f} us  sum {t.sum():.1f} us; blocking call {1e6 * np.median(lat):.0f} us (min {1e6 * min(lat):.0f})", flush=True)
    det.close()

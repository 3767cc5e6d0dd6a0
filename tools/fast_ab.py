"""A/B timing + bit-identity of the FAST stage variants (python tools/fast_ab.py)
fast_var bit 0 = queued score kernel, bit 1 = fused scan + compaction + orientation, bit 2 = 64x32 tiles,
bit 3 = 8 CTAs per SM, bit 4 = pre-test in 16-bit lanes (4 positions per thread)"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import coloc_b200 as clb
from coloc_b200 import synth
clb.set_timing(True)
variants = [int(v) for v in sys.argv[1].split(",")] if len(sys.argv) > 1 else [0, 1, 2, 3, 7, 23, 31, 55]
for (w, h) in ((1280, 720), (3840, 2160), (641, 479), (100, 37)):
    img = synth.config_frame(w, h, 0) if w >= 640 else synth.textured_frame(w, h, 5)
    det = clb.Detector(w, h, maxkp=600000)
    ref = None
    for rep in range(2):
        for var in variants:
            clb.set_param("fast_var", var)
            ts = []
            for _ in range(8):
                k, d = det.detect_and_describe(img)
                ts.append([clb.last_kernel_ms(x) for x in ("resize", "fast", "describe")])
            t = np.min(np.array(ts[1:]), axis=0)
            if ref is None:
                ref = (k.copy(), d.copy())
            same = np.array_equal(ref[0].view(np.uint8), k.view(np.uint8)) and np.array_equal(ref[1], d)
            print(f"{w}x{h} fast_var={var}: {len(k)} kp  fast {t[1]*1e3:.1f} us  (resize {t[0]*1e3:.1f}, describe {t[2]*1e3:.1f})  same={same}", flush=True)
    det.close()
clb.set_param("fast_var", 55)

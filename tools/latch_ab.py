"""A/B timing of the describe kernel: texture vs LDG gather x kernel variants (python tools/latch_ab.py [w h])
latch_var bit 0 = block-shaped gather + lane transpose, bit 1 = schedule table in shared memory,
bit 2 = single gather path (clamp min/max for every keypoint)"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import coloc_b200 as clb
from coloc_b200 import synth
w, h = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1280, 720)
seed = int(sys.argv[3]) if len(sys.argv) > 3 else 1000
img = synth.textured_frame(w, h, seed)
det = clb.Detector(w, h, maxkp=600000)
clb.set_timing(True)
ref = None
for rep in range(2):
    for tex in (0, 1):
        for var in (0, 1, 2, 3, 7):
            clb.set_param("latch_tex", tex)
            clb.set_param("latch_var", var)
            ts = []
            for _ in range(6):
                k, d = det.detect_and_describe(img)
                ts.append([clb.last_kernel_ms(x) for x in ("resize", "fast", "describe")])
            t = np.min(np.array(ts[1:]), axis=0)
            if ref is None:
                ref = d.copy()
            print(f"{w}x{h} tex={tex} var={var}: {len(k)} kp describe {t[2]*1e3:.1f} us  same={np.array_equal(ref, d)}", flush=True)

"""GPU probe: run the REFERENCE CUDALERP kernel on ramp images and save its output, to learn the
texture unit's tex2Dgather footprint rule offline."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle as O
from coloc_b200 import synth
out = {}
for (w, h) in [(640, 480), (1280, 720)]:
    ws, hs, fs = O.level_dims(w, h)
    img = synth.textured_frame(w, h, 6)
    # horizontal and vertical "identity" images: value encodes the source column / row parity pattern
    colid = np.tile((np.arange(w) * 7 % 256).astype(np.uint8), (h, 1))
    rowid = np.tile((np.arange(h) * 7 % 256).astype(np.uint8)[:, None], (1, w))
    out[f"img_{w}"] = img
    for L in range(1, 8):
        out[f"ref_{w}_{L}"] = O.ref_gpu_resize_level(img, fs[L], ws[L], hs[L])
        out[f"col_{w}_{L}"] = O.ref_gpu_resize_level(colid, fs[L], ws[L], hs[L])
        out[f"row_{w}_{L}"] = O.ref_gpu_resize_level(rowid, fs[L], ws[L], hs[L])
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
np.savez_compressed(os.path.join(ROOT, "gpurun_out", "lerp_probe.npz"), **out)
print("saved", len(out))

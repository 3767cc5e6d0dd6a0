import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import coloc_b200 as clb
import oracle as O
from coloc_b200 import synth
w, h = 3840, 2160
img = synth.config_frame(w, h, 0)
okps, odesc = O.detect_describe(img, cap=200000)
opyr = O.pyramid(img)
for trial in range(2):
    det = clb.Detector(w, h, maxkp=200000)
    for rep in range(3):
        kps, desc = det.detect_and_describe(img)
        bad = (desc != odesc).any(axis=1) if len(desc) == len(odesc) else None
        idx = np.nonzero(bad)[0] if bad is not None else []
        print("trial", trial, "rep", rep, len(kps), "kps equal", np.array_equal(kps.view(np.uint8), okps.view(np.uint8)), "bad desc rows", len(idx),
              idx[:8].tolist(), [(int(k["x"]), int(k["y"]), int(k["scale"])) for k in okps[idx[:8]]])
        if len(idx):
            print("   scales of bad rows:", np.bincount(okps["scale"][idx], minlength=8).tolist(), " first/last bad:", int(idx[0]), int(idx[-1]))
            ys = okps["y"][idx]; xs = okps["x"][idx]
            print("   y range", int(ys.min()), int(ys.max()), "x range", int(xs.min()), int(xs.max()))
    pyr = det.pyramid(img)
    print("pyramid diffs", [int((a != b).sum()) for a, b in zip(pyr, opyr)])
    det.close()

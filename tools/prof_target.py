#!/usr/bin/env python3
"""short single-GPU targets for `ncu --set full` captures: `match` (50k x 100k rows), `matchfull` (the bench step: 50k x 2M rows),
`detect` (1280x720 frames), `detect4k` (3840x2160)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import coloc_b200 as clb
from coloc_b200 import synth
what = sys.argv[1] if len(sys.argv) > 1 else "match"
if what in ("match", "matchfull"):
    nq, nt = 50000, (100000 if what == "match" else 2000000)
    T = torch.from_numpy(synth.random_descriptors(nt, 42).view(np.int64)).cuda()
    Q = torch.from_numpy(synth.random_descriptors(nq, 43).view(np.int64)).cuda()
    m = clb.Matcher(maxkp=nt); out = torch.zeros((3, nq), dtype=torch.int32, device="cuda")
    if os.environ.get("CLB_TC_CHUNKS"):          # force the DB chunk count (experiments on the chunk size vs L2)
        clb.set_param("match_balance", 0); clb.set_param("match_tc_chunks", int(os.environ["CLB_TC_CHUNKS"]))
    for _ in range(4):
        m.match_dev(Q, nq, T, nt, 60, out[0], out[1], out[2], stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
else:
    w, h = (3840, 2160) if what == "detect4k" else (1280, 720)
    img = synth.config_frame(w, h, 0)
    det = clb.Detector(w, h, maxkp=400000)
    for _ in range(4):
        k, d = det.detect_and_describe(img)
    print(what, len(k), "keypoints")

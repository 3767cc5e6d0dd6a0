"""GPU probe: bench.stream_extra (frame stream with / without matching), configs 2 and 5."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import coloc_b200 as clb
from coloc_b200 import synth
import bench
for a in ((1280, 720, 256, True), (1280, 720, 256, False), (3840, 2160, 128, True), (3840, 2160, 128, False)):
    r = bench.stream_extra(clb, torch, synth, *a)
    print(a, f"{r['frames_per_s']:.0f} frames/s", f"{r.get('match_cmp_per_s', 0) / 1e12:.2f} T cmp/s", f"matches/frame {r.get('matches_per_frame', 0):.0f}", flush=True)

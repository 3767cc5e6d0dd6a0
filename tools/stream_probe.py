"""GPU probe: bench.stream_extra (frame stream with / without matching) a few times, with the matcher's layout knobs."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import coloc_b200 as clb
from coloc_b200 import synth
import bench
for bal in (-1, 0, -1, 0):
    clb.set_param("match_balance", bal)
    r = bench.stream_extra(clb, torch, synth, 1280, 720, 256, True)
    print(f"match_balance={bal}: {r['frames_per_s']:.0f} frames/s  ({r['match_cmp_per_s'] / 1e12:.2f} T cmp/s)  last balanced={clb.get_param('last_match_balanced')}", flush=True)
r = bench.stream_extra(clb, torch, synth, 1280, 720, 256, False)
print(f"no match: {r['frames_per_s']:.0f} frames/s")

"""stream_extra of bench.py alone (world 1) on the visible GPU: used to tell launcher / NCCL effects from host contention
(run two copies at once on two GPUs).  python tools/stream_pair_probe.py tag"""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
import coloc_b200 as clb
from coloc_b200 import synth
dev = torch.device("cuda", 0); torch.cuda.set_device(0)
out = {}
for key, a in (("c2s", (1280, 720, 256, False)), ("c2m", (1280, 720, 256, True)), ("c5m", (3840, 2160, 128, True)), ("c2s_again", (1280, 720, 256, False))):
    r = bench.stream_extra(clb, torch, None, synth, *a, rank=0, world=1, dev=dev)
    out[key] = round(r["frames_per_s"], 1)
print(sys.argv[1] if len(sys.argv) > 1 else "", json.dumps(out), flush=True)

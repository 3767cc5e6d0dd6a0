"""pair-sized matcher timing vs the DB chunk count (python tools/tc_small.py nq nt chunks...)"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import coloc_b200 as clb
from coloc_b200 import synth
nq, nt = int(sys.argv[1]), int(sys.argv[2])
T = torch.from_numpy(synth.random_descriptors(nt, 42).view(np.int64)).cuda()
Q = torch.from_numpy(synth.random_descriptors(nq, 43).view(np.int64)).cuda()
m = clb.Matcher(maxkp=nt); out = torch.zeros((3, nq), dtype=torch.int32, device="cuda")
clb.set_timing(True)
for ch in [int(x) for x in sys.argv[3:]]:
    # chunks = -1: balanced job list; otherwise uniform chunks (0 = planner's choice)
    clb.set_param("match_balance", 1 if ch == -1 else (-1 if ch == -2 else 0))     # -2: let the time model choose the layout
    clb.set_param("match_tc_chunks", max(ch, 0))
    ts = []
    for _ in range(6):
        m.match_dev(Q, nq, T, nt, 40, out[0], out[1], out[2], stream=torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        ts.append((clb.last_kernel_ms("match"), clb.last_kernel_ms("expand"), clb.last_kernel_ms("merge")))
    t = np.min(np.array(ts[1:]), axis=0)
    print(f"chunks={ch} (balanced={clb.get_param('last_match_balanced')}): match {t[0]*1e3:.1f} us expand(last) {t[1]*1e3:.1f} us merge {t[2]*1e3:.1f} us -> {nq*nt/t[0]/1e9:.2f} T cmp/s", flush=True)

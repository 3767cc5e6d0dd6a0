"""GPU probe: one frame-pair-sized 2-NN call (device pointers) at a few sizes: event time per call; run it under
`ncu --metrics gpu__time_duration.sum` to get the launches behind a call."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import coloc_b200 as clb
from coloc_b200 import synth
sizes = [(int(a), int(b)) for a, b in (s.split("x") for s in sys.argv[1].split(","))] if len(sys.argv) > 1 else [(5000, 5000), (20000, 20000), (50000, 50000)]
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 50
st = torch.cuda.current_stream().cuda_stream
for nq, nt in sizes:
    T = torch.from_numpy(synth.random_descriptors(nt, 42).view(np.int64)).cuda()
    Q = torch.from_numpy(synth.random_descriptors(nq, 43).view(np.int64)).cuda()
    m = clb.Matcher(maxkp=max(nq, nt) + 16); out = torch.zeros(nq, dtype=torch.int32, device="cuda")
    for _ in range(3):
        m.match_dev(Q, nq, T, nt, 40, out, stream=st)
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        m.match_dev(Q, nq, T, nt, 40, out, stream=st)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print(f"{nq} x {nt}: {ms * 1e3:.1f} us/call  {nq * nt / ms / 1e9:.2f} T cmp/s", flush=True)
    m.close()

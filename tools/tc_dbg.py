"""timing experiments on the tensor-core matcher (debug modes give garbage results; times only)"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import coloc_b200 as clb
from coloc_b200 import synth
nq, nt = 50000, 400000
T = torch.from_numpy(synth.random_descriptors(nt, 42).view(np.int64)).cuda()
Q = torch.from_numpy(synth.random_descriptors(nq, 43).view(np.int64)).cuda()
m = clb.Matcher(maxkp=nt); out = torch.zeros((3, nq), dtype=torch.int32, device="cuda")
clb.set_timing(True)
clb.set_param("match_tc", 1)
for spec in sys.argv[1:]:
    kv = dict(x.split("=") for x in spec.split(","))
    for k, v in kv.items():
        clb.set_param(k, int(v))
    ts = []
    for _ in range(5):
        m.match_dev(Q, nq, T, nt, 60, out[0], out[1], out[2], stream=torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        ts.append(clb.last_kernel_ms("match"))
    t = min(ts[1:])
    print(f"{spec}: {t:.3f} ms  {nq * nt / t / 1e9:.1f} T cmp/s... ({nq*nt/t/1e6:.0f} G cmp/s)", flush=True)

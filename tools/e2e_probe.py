"""GPU probe: end-to-end C4 step (pinned host DB -> match list on the host) through clb_match_partial_hostdb for several slice sizes.
python tools/e2e_probe.py"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import coloc_b200 as clb
from coloc_b200 import synth

nq, nt = 50000, 2000000
T = synth.random_descriptors(nt, 42); Q, _ = synth.planted_queries(T, nq, 43)
Tp = torch.from_numpy(T.view(np.int64)).pin_memory(); Qp = torch.from_numpy(Q.view(np.int64)).pin_memory()
m = clb.Matcher(maxkp=nq + 16)
part = torch.zeros((nq, 4), dtype=torch.int32, device="cuda"); out = torch.zeros((3, nq), dtype=torch.int32, device="cuda")
st = torch.cuda.current_stream().cuda_stream

def step():
    q = Qp.to("cuda", non_blocking=True)
    m.match_partial_hostdb(q, nq, Tp, nt, 0, part, stream=st)
    m.merge_partials_dev(part, 1, nq, 60, out[0], out[1], out[2], stream=st)
    return out[0].cpu()

ref = None
for rows in (2097152, 524288, 262144, 131072, 65536, 32768):
    clb.set_param("match_slice_rows", rows)
    for _ in range(2):
        r = step()
    if ref is None:
        ref = r.clone()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        r = step()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"first slice {rows:8d} rows: {ms:.2f} ms/step  {nq * nt / ms / 1e9:.2f} T cmp/s  same={bool(torch.equal(r, ref))}", flush=True)

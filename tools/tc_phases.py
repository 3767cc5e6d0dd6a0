"""GPU probe: phase clocks of one matcher CTA (match_tc_dbg = 7) for short sweeps -- one wave of 148 query blocks
against 8 / 32 / 84 / 160 DB tiles -- and the event time of the same calls."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import coloc_b200 as clb
from coloc_b200 import synth
nq = 37888
clb.set_timing(True)
for nt in (1024, 4096, 10752, 20480):
    T = torch.from_numpy(synth.random_descriptors(nt, 42).view(np.int64)).cuda()
    Q = torch.from_numpy(synth.random_descriptors(nq, 43).view(np.int64)).cuda()
    m = clb.Matcher(maxkp=nt); out = torch.zeros((3, nq), dtype=torch.int32, device="cuda")
    clb.set_param("match_balance", 0); clb.set_param("match_tc_chunks", 1)
    st = torch.cuda.current_stream().cuda_stream
    ts = []
    for _ in range(5):
        m.match_dev(Q, nq, T, nt, 40, out[0], out[1], out[2], stream=st)
        torch.cuda.synchronize(); ts.append(clb.last_kernel_ms("match"))
    print(f"{nq} x {nt} ({nt // 128} tiles per CTA): match kernel {min(ts[1:]) * 1e3:.1f} us", flush=True)
    clb.set_param("match_tc_dbg", 7)
    m.match_dev(Q, nq, T, nt, 40, out[0], out[1], out[2], stream=st)
    torch.cuda.synchronize()
    clb.set_param("match_tc_dbg", 0)
    m.close()
clb.set_param("match_balance", -1); clb.set_param("match_tc_chunks", 0)

"""GPU probe (lab library: CLB_LIB=lab): per-tile clocks of one matcher CTA over the first tiles of a sweep -- when the TMA warp had
issued tile j's operands, when the MMA warp had issued its MMAs, when the accumulator reached epilogue warp 0 and when that warp was
done with it.  Shows who is late while a sweep starts up (tc_cta_cost_us charges ~0.29 us per tile over the first ~50 tiles)."""
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import coloc_b200 as clb
from coloc_b200 import synth
assert clb.is_lab(), "run with CLB_LIB=lab"
nq, nt = 37888, int(sys.argv[1]) if len(sys.argv) > 1 else 20480
T = torch.from_numpy(synth.random_descriptors(nt, 42).view(np.int64)).cuda()
Q = torch.from_numpy(synth.random_descriptors(nq, 43).view(np.int64)).cuda()
m = clb.Matcher(maxkp=nt); out = torch.zeros((3, nq), dtype=torch.int32, device="cuda")
clb.set_param("match_balance", 0); clb.set_param("match_tc_chunks", 1)
st = torch.cuda.current_stream().cuda_stream
for _ in range(3):
    m.match_dev(Q, nq, T, nt, 40, out[0], out[1], out[2], stream=st)
torch.cuda.synchronize()
clb.set_param("match_tc_dbg", 7)
m.match_dev(Q, nq, T, nt, 40, out[0], out[1], out[2], stream=st)
torch.cuda.synchronize()
clb.set_param("match_tc_dbg", 0)
buf = np.zeros((6, 256), np.int64)
rc = clb.lib().clb_lab_tile_clocks(buf.ctypes.data_as(C.c_void_p)); assert rc == 0
n = min(256, nt // 128)
t0 = buf[0, 0]
names = ("tma issued", "mma issued", "acc at epilogue", "epilogue done")
print("tile | " + " | ".join(f"{x:>16s}" for x in names) + " | epilogue busy | acc-to-acc")
for j in list(range(0, min(n, 24))) + list(range(24, n, 8)):
    r = [int(buf[k, j] - t0) for k in range(4)]
    print(f"{j:4d} | " + " | ".join(f"{x:16d}" for x in r) + f" | {int(buf[3, j] - buf[2, j]):13d} | {int(buf[2, j] - buf[2, j - 1]) if j else 0:10d}")
d = np.diff(buf[2, :n])
for a, b in ((0, 8), (8, 16), (16, 32), (32, 64), (64, 128), (128, 160)):
    if b <= n:
        print(f"tiles {a:3d}-{b:3d}: {d[a:b - 1].mean():7.0f} cycles per tile (accumulator to accumulator), epilogue busy {np.mean(buf[3, a:b] - buf[2, a:b]):6.0f}"
              f" = TMEM loads {np.mean(buf[4, a:b] - buf[2, a:b]):5.0f} + columns 0-31 {np.mean(buf[5, a:b] - buf[4, a:b]):5.0f} + columns 32-63 {np.mean(buf[3, a:b] - buf[5, a:b]):5.0f}")

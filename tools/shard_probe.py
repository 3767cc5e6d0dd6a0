"""one GPU, the 8-GPU shard shape: 50k queries x 250k host-resident rows through clb_sharded_match_map (rank mode, world 1),
first-slice size varied; median call time + event timeline.  python tools/shard_probe.py [rows]"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
import coloc_b200 as clb
NQ = bench.NQ; nt = int(sys.argv[1]) if len(sys.argv) > 1 else 250000
T_host = torch.from_numpy(bench.make_db_shard(0, nt).view(np.int64)).pin_memory()
Qn, _ = bench.make_queries(NQ, bench.NT)
Q_host = torch.from_numpy(Qn.view(np.int64)).pin_memory()
dT = T_host.cuda(); dQ = Q_host.cuda()
sm = clb.ShardedMatcher(device=0, rank=0, nranks=1, thresh=60, maxkp=nt + 16)
out = torch.zeros((3, NQ), dtype=torch.int32, device="cuda"); st = torch.cuda.current_stream().cuda_stream
mh = torch.zeros(NQ, dtype=torch.int32).pin_memory().numpy()
def med(step, n=20):
    for _ in range(4): step()
    torch.cuda.synchronize(); ts = []
    for _ in range(n):
        t0 = time.perf_counter(); step(); torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
    return 1e3 * float(np.median(ts))
sm.set_map_shard(dT, 0, nt, where=1)
print(f"resident {nt} rows: {med(lambda: sm.match_map_dev(dQ, NQ, 60, out[0], out[1], out[2], stream=st)):.3f} ms", flush=True)
ref = out[0].cpu().numpy().copy()
sm.set_map_shard(T_host, 0, nt, where=2)
for rows in (1 << 30, 131072, 65536, 32768, 16384, 8192):
    clb.set_param("match_slice_rows", rows)
    ms = med(lambda: sm.match_map(Q_host, 60, out=mh))
    ok = np.array_equal(mh, ref)
    clb.set_timing(2); clb.timeline(); sm.match_map(Q_host, 60, out=mh); tl = clb.timeline(); clb.set_timing(0)
    print(f"first slice {rows}: {ms:.3f} ms same={ok}  " + " ".join(f"{a}:{b:.3f}" for a, b in tl), flush=True)
clb.set_param("match_slice_rows", 131072)

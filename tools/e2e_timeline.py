"""event timeline of one sharded C4 step (resident and host-DB e2e) on rank 0
(torchrun --nproc-per-node N tools/e2e_timeline.py; N = 1 works without torchrun)"""
import os, sys, time, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import bench
import coloc_b200 as clb
from coloc_b200 import shard
rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); lr = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(lr)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
dev = torch.device("cuda", lr)
NQ, NT, THR = bench.NQ, bench.NT, bench.THRESH
lo, hi = shard.shard_range(NT, rank, world)
T_host = torch.from_numpy(bench.make_db_shard(lo, hi).view(np.int64)).pin_memory()
Qn, _ = bench.make_queries(NQ, NT)
Q_host = torch.from_numpy(Qn.view(np.int64)).pin_memory()
dT = T_host.to(dev); dQ = Q_host.to(dev)
uid = [clb.ShardedMatcher.unique_id() if rank == 0 else None]
if world > 1:
    dist.broadcast_object_list(uid, src=0)
sm = clb.ShardedMatcher(device=lr, rank=rank, nranks=world, uid=uid[0], thresh=THR, maxkp=max(hi - lo, NQ) + 16)
out = torch.zeros((3, NQ), dtype=torch.int32, device=dev)
st = torch.cuda.current_stream().cuda_stream
mh = torch.zeros(NQ, dtype=torch.int32).pin_memory().numpy()
res = {"ranks": world, "rows_per_rank": hi - lo}
for mode in ("resident", "e2e"):
    if mode == "resident":
        sm.set_map_shard(dT, lo, NT, where=1)
        step = lambda: sm.match_map_dev(dQ, NQ, THR, out[0], out[1], out[2], stream=st)
    else:
        sm.set_map_shard(T_host, lo, NT, where=2)
        step = lambda: sm.match_map(Q_host, THR, out=mh)
    for _ in range(5):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ts = []
    for _ in range(20):
        t0 = time.perf_counter(); step(); torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
    res[mode + "_ms_median"] = 1e3 * float(np.median(ts))
    if world > 1:
        dist.barrier()
    clb.set_timing(2); clb.timeline()
    step(); torch.cuda.synchronize()
    res[mode + "_timeline_ms"] = clb.timeline()
    clb.set_timing(0)
if rank == 0:
    print(json.dumps(res))
sm.close()
if world > 1:
    dist.destroy_process_group()

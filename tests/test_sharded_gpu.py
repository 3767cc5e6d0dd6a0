"""GPU tests of the sharded map matcher in the C ABI (clb_sharded_*): GPUMatcher::setMapData + matchFeaturesWithMap
(GPUMatcher.hpp:110-117, 252-271) with the map rows cut over several GPUs, against the oracle's single scan -- ties across shard
borders included.  The device-list and NCCL-rank tests need 2+ GPUs and skip below that (run them with `gpurun --gpus 2`)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from coloc_b200 import synth   # noqa: E402

pytestmark = pytest.mark.gpu


def _problem(nt=30011, nq=1500, seed=5):
    T = synth.random_descriptors(nt, seed); Q, _ = synth.planted_queries(T, nq, seed + 1)
    # the same row in the first, a middle and the last shard (any shard count up to 8): the lowest global index must win
    T[nt - 1] = T[3]; T[nt // 2 + 1] = T[3]; Q[0] = T[3]
    # a duplicate on both sides of every shard border of 2, 3, 4 and 8 shards
    for world in (2, 3, 4, 8):
        base, rem = divmod(nt, world)
        for r in range(1, world):
            lo = r * base + min(r, rem)
            T[lo] = T[lo - 1]
    Q[1:9] = T[[nt // 2 - 1, nt // 2, nt // 3 - 1, nt // 3, nt // 4, nt // 8, nt - 2, 7]]
    return T, Q


def _check(got, ref):
    for name, a, b in zip(("match", "best", "second"), got, ref):
        assert np.array_equal(a, b), (name, int((a != b).sum()))


def _ngpu():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.parametrize("tc", [-1, 1, 0])
def test_sharded_one_device(clb, oracle, tc):
    """one shard: the plumbing (query staging, resident and host-resident map, thresholds) on a single GPU"""
    T, Q = _problem()
    clb.set_param("match_tc", tc)
    try:
        sm = clb.ShardedMatcher(devices=[0], thresh=60, maxkp=4096)
        assert sm.n == 1 and sm.shard_range(len(T), 0) == (0, len(T))
        sm.set_map(T)
        _check(sm.match_map(Q, want_dist=True), oracle.match_2nn(Q, T, 60))          # threshold from opts
        _check(sm.match_map(Q[:77], 40, want_dist=True), oracle.match_2nn(Q[:77], T, 40))
        assert len(sm.match_map(Q[:0])) == 0
        sm.set_map(T, host_resident=True)                                           # rows stream from host memory per call
        _check(sm.match_map(Q, 60, want_dist=True), oracle.match_2nn(Q, T, 60))
        sm.close()
        one = clb.ShardedMatcher(device=0, rank=0, nranks=1, thresh=60, maxkp=4096)  # rank mode, world of one: no NCCL needed
        one.set_map_shard(T, 0, len(T), where=0)
        _check(one.match_map(Q, 60, want_dist=True), oracle.match_2nn(Q, T, 60))
        one.close()
    finally:
        clb.set_param("match_tc", -1)


def test_sharded_bad_arguments(clb):
    with pytest.raises(clb.ClbError):
        clb.ShardedMatcher(devices=[0, 0])
    with pytest.raises(clb.ClbError):
        clb.ShardedMatcher(devices=[_ngpu() + 3])
    with pytest.raises(clb.ClbError):
        clb.ShardedMatcher(device=0, rank=2, nranks=2, uid=b"\0" * 128)


@pytest.mark.parametrize("tc", [-1, 1])
def test_sharded_device_list(clb, oracle, tc):
    """one process, every visible device: sweeps run concurrently, partials reach device 0 by peer copies, merge there"""
    n = _ngpu()
    if n < 2:
        pytest.skip("needs 2+ GPUs")
    T, Q = _problem()
    ref = oracle.match_2nn(Q, T, 60)
    clb.set_param("match_tc", tc)
    try:
        for ndev in sorted({2, n}):
            sm = clb.ShardedMatcher(devices=list(range(ndev)), thresh=60, maxkp=4096)
            sm.set_map(T)
            for _ in range(2):
                _check(sm.match_map(Q, want_dist=True), ref)
            sm.set_map(T, host_resident=True)
            _check(sm.match_map(Q, want_dist=True), ref)
            sm.set_map(T[:5], host_resident=False)                                  # fewer rows than devices: empty shards
            _check(sm.match_map(Q[:50], 60, want_dist=True), oracle.match_2nn(Q[:50], T[:5], 60))
            sm.close()
    finally:
        clb.set_param("match_tc", -1)


def _rank_worker(rank, world, uid_path, out_q):
    try:
        import torch
        torch.cuda.set_device(rank)
        import coloc_b200 as clb
        import oracle as O
        T, Q = _problem()
        if rank == 0:
            uid = clb.ShardedMatcher.unique_id()
            with open(uid_path + ".tmp", "wb") as f:
                f.write(uid)
            os.replace(uid_path + ".tmp", uid_path)
        else:
            import time
            for _ in range(600):
                if os.path.exists(uid_path):
                    break
                time.sleep(0.05)
            uid = open(uid_path, "rb").read()
        sm = clb.ShardedMatcher(device=rank, rank=rank, nranks=world, uid=uid, thresh=60, maxkp=4096)
        lo, hi = sm.shard_range(len(T), rank)
        ref = O.match_2nn(Q, T, 60)
        ok = True
        for where in (0, 1, 2):
            rows = T[lo:hi]
            if where == 1:
                rows = torch.from_numpy(rows.view(np.int64).copy()).cuda()
            elif where == 2:
                rows = torch.from_numpy(rows.view(np.int64).copy()).pin_memory()
            sm.set_map_shard(rows, lo, len(T), where=where)
            got = sm.match_map(Q, want_dist=True)
            ok = ok and all(np.array_equal(a, b) for a, b in zip(got, ref))
        # device entry on torch's stream
        dQ = torch.from_numpy(Q.view(np.int64)).cuda()
        out = torch.zeros((3, len(Q)), dtype=torch.int32, device="cuda")
        sm.set_map(T)                                                               # whole map: every rank keeps its own rows
        sm.match_map_dev(dQ, len(Q), 60, out[0], out[1], out[2], stream=torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        ok = ok and all(np.array_equal(out[i].cpu().numpy(), ref[i]) for i in range(3))
        sm.close()
        out_q.put((rank, ok, ""))
    except Exception as e:      # noqa: BLE001
        import traceback
        out_q.put((rank, False, traceback.format_exc() + repr(e)))


def test_sharded_ranks_nccl(clb, tmp_path):
    """one process per GPU: the exchange is one ncclAllGather inside clb_sharded_match_map*; every rank gets the oracle's list"""
    n = _ngpu()
    if n < 2:
        pytest.skip("needs 2+ GPUs")
    import torch.multiprocessing as mp
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    uid_path = str(tmp_path / "nccl_uid.bin")
    procs = [ctx.Process(target=_rank_worker, args=(r, world, uid_path, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=600) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    for rank, ok, err in res:
        assert ok, (rank, err)

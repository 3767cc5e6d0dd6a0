"""world_size-2 (and 3) gloo tests of the host-side sharding logic on CPU: shard ranges, the pair/row plan,
the all-gather layout of partials, and that (oracle partials per shard) + gather + merge rule == the
oracle's single scan, ties included.  The device merge kernel itself is covered by -m gpu tests."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from coloc_b200 import shard, synth   # noqa: E402


def test_shard_range_covers():
    for n in (0, 1, 7, 8, 9, 2000000):
        for world in (1, 2, 3, 8):
            spans = [shard.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def test_pair_plan_covers_every_row_once():
    pairs = shard.ordered_pairs(4)
    assert len(pairs) == 12 and (0, 0) not in pairs
    counts = {0: 5, 1: 0, 2: 17, 3: 8}
    for world in (1, 2, 4, 8):
        seen = []
        for r in range(world):
            for p, a, b in shard.pair_row_plan(pairs, counts, r, world):
                assert 0 <= a < b <= counts[pairs[p][0]]
                seen += [(p, i) for i in range(a, b)]
        want = [(p, i) for p, (q, _) in enumerate(pairs) for i in range(counts[q])]
        assert seen == want


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    try:
        os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        import oracle as O
        from _helpers import merge_partials_host
        O.set_threads(1)
        nq, nt = 64, 1001
        T = synth.random_descriptors(nt, 5); Q, _ = synth.planted_queries(T, nq, 6)
        T[700] = T[3]; T[999] = T[3]; Q[0] = T[3]            # tie across shards: lowest global index must win
        lo, hi = shard.shard_range(nt, rank, world)
        m, b, s = O.match_2nn(Q, T[lo:hi], -1)                # threshold -1: m = local best index
        part = np.zeros(nq, dtype=[("best_d", "<i4"), ("best_i", "<i4"), ("second_d", "<i4"), ("_pad", "<i4")])
        part["best_d"] = b; part["second_d"] = s; part["best_i"] = np.where(m >= 0, m + lo, -1)
        allp = shard.gather_partials(part).numpy()            # [world, nq, 4] int32
        assert allp.shape == (world, nq, 4)
        assert np.array_equal(allp[rank].reshape(-1), part.view(np.int32).reshape(-1))
        parts = allp.reshape(world, nq * 4).view(part.dtype).reshape(world, nq)
        got = merge_partials_host(parts, 40)
        ref = O.match_2nn(Q, T, 40)
        ok = all(np.array_equal(a, b) for a, b in zip(got, ref))
        # every rank ends with the same list (all-gather, not gather)
        t = torch.from_numpy(got[0].copy()); ref0 = t.clone(); dist.broadcast(ref0, 0)
        ok = ok and bool((t == ref0).all())
        q.put((rank, ok, None))
        dist.destroy_process_group()
    except Exception as e:   # pragma: no cover
        q.put((rank, False, repr(e)))


@pytest.mark.parametrize("world", [2, 3])
def test_db_sharding_gloo(world):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok, _ in res), res

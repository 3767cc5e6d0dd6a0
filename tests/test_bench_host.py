"""Host-side logic of bench.py that the numbers depend on and that needs no GPU: the synthetic C4 workload must be the same whatever
the rank count (every rank builds only its DB shard, blockwise), the planted queries must point at the rows they were made from, and
the rank -> core-block assignment must be disjoint and stay inside the GPU's own CPU set where NVML reports one."""
import importlib.util
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def bench():
    spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(ROOT, "bench.py"))
    mod = importlib.util.module_from_spec(spec)
    argv = sys.argv
    sys.argv = ["bench.py"]
    try:
        spec.loader.exec_module(mod)
    finally:
        sys.argv = argv
    return mod


def test_db_shards_tile_the_same_database(bench):
    nt = 600001                                             # crosses two generation blocks of 250 000 rows, ragged end
    whole = bench.make_db_shard(0, nt)
    assert whole.shape == (nt, 8) and whole.dtype == np.uint64
    for world in (2, 3, 8):
        base, rem = divmod(nt, world)
        parts = []
        for r in range(world):
            lo = r * base + min(r, rem)
            parts.append(bench.make_db_shard(lo, lo + base + (1 if r < rem else 0)))
        assert np.array_equal(np.concatenate(parts), whole), world
    assert np.array_equal(bench.make_db_shard(249990, 250010), whole[249990:250010])          # a shard border inside a block border


def test_queries_are_planted_on_the_rows_they_name(bench, oracle):
    nt, nq = 500000, 400
    Q, src = bench.make_queries(nq, nt)
    T = bench.make_db_shard(0, nt)
    planted = src >= 0
    assert abs(planted.mean() - 0.8) < 0.01 and Q.shape == (nq, 8)
    d = np.array([sum(bin(int(w)).count("1") for w in (Q[i] ^ T[src[i]])) for i in np.nonzero(planted)[0]])
    assert 20 < d.mean() < 80 and d.max() < 110            # each bit flipped with probability 0.10: ~51 of 512
    # the oracle finds them (same generator as the bench line's planted_match_recall)
    m, best, second = oracle.match_2nn(Q[:64], T, 60)
    want = src[:64]
    assert np.array_equal(m[want >= 0], want[want >= 0])


def test_rank_core_blocks(bench, monkeypatch):
    if not hasattr(os, "sched_getaffinity"):
        pytest.skip("no sched_setaffinity on this platform")
    before = os.sched_getaffinity(0)
    cores = sorted(before)
    if len(cores) < 4:
        pytest.skip("needs 4+ cores")
    try:
        # no topology information: plain disjoint blocks
        monkeypatch.setattr(bench, "_nvml_device_cpus", lambda world: None)
        seen = []
        for r in range(4):
            os.sched_setaffinity(0, before)
            bench.pin_rank_to_cores(r, 4)
            mine = os.sched_getaffinity(0)
            assert bench._AFFINITY["mode"] == "blocks" and len(mine) == len(cores) // 4
            seen.append(mine)
        assert all(a.isdisjoint(b) for i, a in enumerate(seen) for b in seen[i + 1:])
        # a VM that reports the same set for every GPU (the 8-GPU box of this round): still plain blocks
        monkeypatch.setattr(bench, "_nvml_device_cpus", lambda world: [set(cores)] * world)
        os.sched_setaffinity(0, before)
        bench.pin_rank_to_cores(1, 4)
        assert bench._AFFINITY["mode"] == "blocks"
        # two sockets: ranks 0, 1 on the first half of the cores, ranks 2, 3 on the second
        half = len(cores) // 2
        monkeypatch.setattr(bench, "_nvml_device_cpus", lambda world: [set(cores[:half]) if r < world // 2 else set(cores[half:]) for r in range(world)])
        seen = []
        for r in range(4):
            os.sched_setaffinity(0, before)
            bench.pin_rank_to_cores(r, 4)
            mine = os.sched_getaffinity(0)
            assert bench._AFFINITY["mode"] == "nvml-numa"
            assert mine <= (set(cores[:half]) if r < 2 else set(cores[half:]))
            seen.append(mine)
        assert all(a.isdisjoint(b) for i, a in enumerate(seen) for b in seen[i + 1:])
    finally:
        os.sched_setaffinity(0, before)

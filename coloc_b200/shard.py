"""Multi-GPU sharding of the matching configs, one process per GPU.

  * DB-row sharding (C4, map relocalisation): rank g owns rows [lo, hi); every rank holds all
    queries, computes the per-query partial (best_d, best_i_global, second_d) on its rows, ONE
    all-gather of nQ x 16 B per rank moves the partials (the only exchange step of the path), and
    a merge kernel folds them in ascending rank order -- which preserves first-min-wins -- and
    applies the margin test.  Bit-exact with the single-GPU scan by construction.
  * query-row sharding (C3, 12 ordered pairs): the flattened (pair, query-row) list is cut into
    contiguous slices; no exchange inside the data path, only a gather of the int32 match lists
    (-1 = none: the reference's g_match format, CUDAK2NN.cu:75).

The sharded matcher itself lives in the C ABI (clb_sharded_*, coloc_b200/csrc/api_sharded.cu; Python binding
coloc_b200.ShardedMatcher): sweep + ncclAllGather + merge inside the library.  This module keeps the pure index helpers
(shard ranges, the pair / row plan) that bench.py and the CPU (gloo) tests share.
"""
import numpy as np


def shard_range(n, rank, world):
    """contiguous, balanced [lo, hi) slice of n rows for `rank` of `world` (first n % world ranks get one more)."""
    base, rem = divmod(int(n), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def ordered_pairs(n_frames):
    """all ordered inter-vehicle pairs (i != j); BASELINE config 3.  (The reference itself enumerates the
    unordered half, Pair_Builder.hpp:25-33; the ordered list is that plus the swapped pairs.)"""
    return [(i, j) for i in range(n_frames) for j in range(n_frames) if i != j]


def pair_row_plan(pairs, counts, rank, world):
    """query-row sharding of a pair list: flatten (pair, query row) and give `rank` a contiguous slice.
    Returns [(pair_index, q_lo, q_hi)] covering this rank's rows, in order."""
    sizes = [counts[q] for q, _ in pairs]
    total = int(sum(sizes))
    lo, hi = shard_range(total, rank, world)
    out, off = [], 0
    for p, n in enumerate(sizes):
        a, b = max(lo, off), min(hi, off + n)
        if a < b:
            out.append((p, a - off, b - off))
        off += n
    return out


def gather_partials(local_parts, group=None):
    """all-gather of per-rank partial arrays ([nQ] PARTIAL_DTYPE numpy or [nQ,4] int32 tensor) -> [world, nQ, 4]
    int32 tensor.  Works on gloo (CPU tensors) and NCCL (CUDA tensors)."""
    import torch
    import torch.distributed as dist
    if isinstance(local_parts, np.ndarray):
        local_parts = torch.from_numpy(np.ascontiguousarray(local_parts).view(np.int32).reshape(-1, 4))
    world = dist.get_world_size(group)
    out = torch.empty((world,) + tuple(local_parts.shape), dtype=local_parts.dtype, device=local_parts.device)
    dist.all_gather_into_tensor(out.view(-1), local_parts.contiguous().view(-1), group=group)
    return out

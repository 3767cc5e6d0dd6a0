"""Multi-GPU sharding of the matching configs, one process per GPU.

  * DB-row sharding (C4, map relocalisation): rank g owns rows [lo, hi); every rank holds all
    queries, computes the per-query partial (best_d, best_i_global, second_d) on its rows, ONE
    all-gather of nQ x 16 B per rank moves the partials (the only exchange step of the path), and
    a merge kernel folds them in ascending rank order -- which preserves first-min-wins -- and
    applies the margin test.  Bit-exact with the single-GPU scan by construction.
  * query-row sharding (C3, 12 ordered pairs): the flattened (pair, query-row) list is cut into
    contiguous slices; no exchange inside the data path, only a gather of the int32 match lists
    (-1 = none: the reference's g_match format, CUDAK2NN.cu:75).

torch / torch.distributed are plumbing here (device buffers, NCCL or gloo collectives); all
arithmetic is in libcoloc_b200.so.  The pure index helpers at the top are CPU-testable (gloo).
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


class ShardedMatcher:
    """DB-row-sharded 2-NN over a torch.distributed process group (NCCL on GPUs).

    set_shard() uploads this rank's DB rows once (GPUMatcher::setMapData analogue); match() is the
    per-frame call: H2D of the queries, local partial, all-gather, device merge, D2H of match[]."""

    def __init__(self, matcher, group=None):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.m = matcher
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.dT = None
        self.n_local = 0
        self.t_base = 0

    def set_shard_dev(self, dT_rows, t_base):
        """dT_rows: int64 CUDA tensor [n_local, 8] (device-resident shard), t_base: global index of row 0."""
        self.dT, self.n_local, self.t_base = dT_rows, int(dT_rows.shape[0]), int(t_base)
        # resident shard: copied and expanded once inside the matcher (GPUMatcher::setMapData analogue), not per step
        self.m.set_map_dev(dT_rows, self.n_local, stream=self.torch.cuda.current_stream().cuda_stream)

    def set_shard(self, T_rows_host, t_base):
        t = self.torch.from_numpy(np.ascontiguousarray(T_rows_host, np.uint64).view(np.int64).reshape(-1, 8))
        self.set_shard_dev(t.cuda(non_blocking=False), t_base)

    def match_dev(self, dQ, nQ, threshold, d_out, parts_buf=None, gather_buf=None):
        """device-resident step.  d_out: int32 CUDA tensor [3, nQ] (match, best, second).  Enqueues on torch's
        current stream; returns nothing."""
        torch, dist = self.torch, self.dist
        stream = torch.cuda.current_stream().cuda_stream   # 0 (legacy default) is mapped to cudaStreamLegacy by the binding
        if self.world == 1:
            self.m.match_map_dev(dQ, nQ, threshold, d_out[0], d_out[1], d_out[2], stream=stream)
            return
        local = parts_buf if parts_buf is not None else torch.empty((nQ, 4), dtype=torch.int32, device=dQ.device)
        allp = gather_buf if gather_buf is not None else torch.empty((self.world, nQ, 4), dtype=torch.int32, device=dQ.device)
        self.m.match_map_partial_dev(dQ, nQ, self.t_base, local, stream=stream)
        dist.all_gather_into_tensor(allp.view(-1), local.view(-1), group=self.group)   # the one exchange step
        self.m.merge_partials_dev(allp, self.world, nQ, threshold, d_out[0], d_out[1], d_out[2], stream=stream)

    def match(self, Q_host_pinned, threshold):
        """end-to-end step from HOST queries (pinned torch int64 [nQ, 8]) to a HOST match list."""
        torch = self.torch
        nQ = int(Q_host_pinned.shape[0])
        dQ = Q_host_pinned.cuda(non_blocking=True)
        out = torch.empty((3, nQ), dtype=torch.int32, device=dQ.device)
        self.match_dev(dQ, nQ, threshold, out)
        return out[0].cpu()


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

"""Seed-sharded expansion over several GPUs of one box.

The path shards by seed: every seed expands independently against a read-only graph, so the graph
is replicated in each GPU's HBM and there is no exchange step inside the traversal.  One process
per GPU (`torch.distributed`).  Collectives, all outside the kernels:
  * all-gather of the per-rank, per-length instance counts (a few int64 per rank) - always;
  * all-reduce style decisions of the metapath ladder: "no rank found a row -> next rung" and the
    global shortest length (reference find_metapath/find_srckind_metapath_neo4j.py:142-154);
  * optionally the rows themselves, gathered to one rank in rank order.  Shards are CONTIGUOUS
    ranges of the sorted seed list, so rank-order concatenation is already ordered by seed.

Nothing here needs a GPU: the same code runs over gloo with CPU tensors (tests/test_parallel.py).
"""
from __future__ import annotations

import numpy as np


def partition_by_work(weights, world: int):
    """Cut range(len(weights)) into `world` contiguous pieces of roughly equal total weight.
    Returns [(lo, hi)] * world; pieces may be empty; every index is in exactly one piece."""
    w = np.asarray(weights, dtype=np.float64)
    n = len(w)
    if n == 0:
        return [(0, 0)] * world
    cum = np.cumsum(np.maximum(w, 0.0) + 1e-9)          # strictly increasing: ties cannot collapse pieces
    cuts = np.searchsorted(cum, cum[-1] * np.arange(1, world) / world, side="left")
    cuts = np.clip(cuts, 0, n)
    edges = np.concatenate([[0], cuts, [n]])
    edges = np.maximum.accumulate(edges)
    return [(int(edges[r]), int(edges[r + 1])) for r in range(world)]


def first_hop_degree(snap, seeds, direction="und", entity_only=True):
    """Predicted work per seed: the length of the adjacency row its first hop scans."""
    seeds = np.asarray(seeds, dtype=np.int64)
    rng = {"out": snap.out_rng, "in": snap.in_rng, "und": snap.und_rng}[direction]
    ok = (seeds >= 0) & (seeds < snap.n_nodes)
    s = np.where(ok, seeds, 0)
    end = rng[s, 1] if entity_only else rng[s + 1, 0]
    return np.where(ok, end - rng[s, 0], 0)


def shard_seeds(snap, seeds, rank: int, world: int, direction="und", entity_only=True, seed_ts=None):
    """This rank's contiguous slice of `seeds` (and of `seed_ts`), balanced by first-hop degree."""
    lo, hi = partition_by_work(1 + first_hop_degree(snap, seeds, direction, entity_only), world)[rank]
    return (seeds[lo:hi], None if seed_ts is None else seed_ts[lo:hi], (lo, hi))


# ------------------------------------------------------------------------------------------ collectives
def _dist():
    import torch.distributed as dist
    return dist


def world_info():
    dist = _dist()
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def allgather_counts(local_counts, device="cpu"):
    """local_counts: per-length instance counts of this rank -> int64 tensor [world, len(local_counts)]."""
    import torch
    rank, world = world_info()
    mine = torch.tensor(list(local_counts), dtype=torch.int64, device=device)
    if world == 1:
        return mine[None, :].clone()
    out = torch.empty(world * mine.numel(), dtype=torch.int64, device=device)      # flat: gloo insists
    _dist().all_gather_into_tensor(out, mine)
    return out.view(world, mine.numel())


def connect_peers(engine):
    """Set up the engine's count exchange over peer memory (rca_peer_*): the CUDA IPC handles of the ranks'
    slot arrays travel through torch.distributed once; afterwards engine.peer_allgather(value) is one small
    kernel on the engine's stream - no NCCL call, nothing on the host - per exchange."""
    rank, world = world_info()
    mine = engine.peer_handle(world)
    if world == 1:
        handles = [mine]
    else:
        handles = [None] * world
        _dist().all_gather_object(handles, mine)
    engine.peer_connect(rank, handles)


def connect_gather(engine, words: int, dst: int = 0):
    """Set up the engine's row gather over peer memory (rca_gather_*): rank `dst` allocates a window of
    `words` int32 words, its CUDA IPC handle travels through torch.distributed, every rank maps it.
    connect_peers(engine) must have been called.  Call again with a larger `words` to replace the window."""
    rank, world = world_info()
    box = [engine.gather_window(words) if rank == dst else None]
    if world > 1:
        _dist().broadcast_object_list(box, src=dst)
    engine.gather_connect(dst, box[0], words)


def gathered_rows(engine, result, level: int, dst: int = 0, offset_words: int = 0):
    """engine.gather_rows as a torch tensor on the destination (zero-copy view of the window), None elsewhere.
    -> (tensor or None, per-rank row counts, this rank's push time in ms)."""
    import torch
    ptr, n, per, ms = engine.gather_rows(result, level, offset_words)
    rank, _ = world_info()
    if rank != dst:
        return None, per, ms
    w = result.row_len[level]
    if n == 0:
        return torch.empty((0, w), dtype=torch.int32, device=f"cuda:{engine.device}"), per, ms
    return torch.as_tensor(_CudaView(ptr, n, w), device=f"cuda:{engine.device}"), per, ms


def ladder_decision(counts_by_rank):
    """From allgather_counts output: (any rows anywhere?, globally shortest length with rows or -1)."""
    tot = counts_by_rank.sum(dim=0)
    nz = (tot > 0).nonzero()
    return bool(len(nz)), (int(nz[0]) if len(nz) else -1)


def gather_rows(rows, counts_for_level, dst: int = 0):
    """Rows of one length from every rank to `dst`, concatenated in rank order.
    rows: this rank's [n, w] int32 tensor (CPU for gloo, CUDA for NCCL); counts_for_level: int64
    tensor [world] (a column of allgather_counts).  Returns the [sum n, w] tensor on dst, None elsewhere."""
    import torch
    dist = _dist()
    rank, world = world_info()
    if world == 1:
        return rows
    w = rows.shape[1]
    if rank == dst:
        total = int(counts_for_level.sum())
        out = torch.empty((total, w), dtype=rows.dtype, device=rows.device)
        offs = torch.cumsum(counts_for_level, 0) - counts_for_level
        ops = []
        for r in range(world):
            n = int(counts_for_level[r])
            if n == 0:
                continue
            view = out[int(offs[r]):int(offs[r]) + n]
            if r == dst:
                view.copy_(rows)
            else:
                ops.append(dist.P2POp(dist.irecv, view, r))
        if ops:
            for req in dist.batch_isend_irecv(ops):
                req.wait()
        return out
    if rows.shape[0]:
        for req in dist.batch_isend_irecv([dist.P2POp(dist.isend, rows.contiguous(), dst)]):
            req.wait()
    return None


class _CudaView:
    """Zero-copy view of the engine's device rows for torch (via __cuda_array_interface__)."""

    def __init__(self, ptr, n, w):
        self.__cuda_array_interface__ = {"shape": (n, w), "typestr": "<i4", "data": (ptr, False), "version": 3,
                                         "strides": None}


def device_rows(result, level):
    """torch tensor over the rows of `level` that still live in the engine's output arena (valid
    until the next expand() on that engine)."""
    import torch
    ptr, n, w = result.device_rows(level)
    if n == 0:
        return torch.empty((0, w), dtype=torch.int32, device=f"cuda:{result.engine.device}")
    return torch.as_tensor(_CudaView(ptr, n, w), device=f"cuda:{result.engine.device}")


class ShardedExpander:
    """expand() over all ranks: every rank expands its contiguous slice of the seeds on its own
    engine, then the counts are all-gathered (and the rows gathered to rank `dst` on request)."""

    def __init__(self, snap, expand_fn, device="cpu"):
        """expand_fn(plan, seeds, seed_ts) -> {length: rows ndarray or tensor} for the local slice.
        With the GPU engine: lambda p, s, t: engine.expand(p, s, t)   (see run_gpu)."""
        self.snap, self.expand_fn, self.device = snap, expand_fn, device

    def run(self, plan_levels: int, plan, seeds, seed_ts=None, direction="und", entity_only=True, gather=False, dst=0):
        import torch
        rank, world = world_info()
        my_seeds, my_ts, (lo, hi) = shard_seeds(self.snap, seeds, rank, world, direction, entity_only, seed_ts)
        levels = self.expand_fn(plan, my_seeds, my_ts)
        local = [int(len(levels[s])) if s in levels else 0 for s in range(plan_levels + 1)]
        counts = allgather_counts(local, self.device)
        any_rows, shortest = ladder_decision(counts)
        out = {"slice": (lo, hi), "counts": counts, "total": int(counts.sum()), "any_rows": any_rows,
               "shortest": shortest, "rows": None}
        if gather:
            rows = {}
            for s in range(plan_levels + 1):
                if int(counts[:, s].sum()) == 0:
                    continue
                mine = levels.get(s)
                if mine is None:
                    width = None
                else:
                    width = mine.shape[1]
                # every rank must agree on the row width even if it holds no row of this length
                wt = torch.tensor([width or 0], dtype=torch.int64, device=self.device)
                if world > 1:
                    _dist().all_reduce(wt, op=_dist().ReduceOp.MAX)
                width = int(wt)
                if mine is None:
                    mine = torch.empty((0, width), dtype=torch.int32, device=self.device)
                elif not isinstance(mine, torch.Tensor):
                    mine = torch.from_numpy(np.ascontiguousarray(mine)).to(self.device)
                rows[s] = gather_rows(mine, counts[:, s], dst)
            out["rows"] = rows
        return out

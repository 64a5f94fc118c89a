"""Worker of tests/test_parallel.py: one rank of a world_size-N gloo job on CPU.  Each rank expands
its slice of the seeds with the CPU oracle (test helper standing in for the per-GPU engine) and the
collectives of k8s_llm_rca_b200.parallel run for real over gloo."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import torch.distributed as dist
    from k8s_llm_rca_b200 import parallel, synth
    from oracle.csr import CsrOracle
    from conftest import sort_rows

    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    snap = synth.generate(2000, 200, seed=31)
    orc = CsrOracle(snap)
    K = snap.kind_index
    pod, node, ev, ns = K("Pod"), K("Node"), K("Event"), K("Namespace")
    seeds = snap.nodes_of_kind(pod)
    kw = dict(dest_kind=node, excl_kinds=[ev, ns], entity_only=True)

    def expand(plan, s, t):
        levels, _ = orc.varlen(s, "und", 3, **kw)
        return levels

    sx = parallel.ShardedExpander(snap, expand, device="cpu")
    out = sx.run(3, None, seeds, gather=True, dst=0)

    # every rank sees every rank's counts; slices are contiguous, disjoint and cover the seeds
    slices = [None] * world
    dist.all_gather_object(slices, out["slice"])
    assert slices[0][0] == 0 and slices[-1][1] == len(seeds)
    assert all(slices[i][1] == slices[i + 1][0] for i in range(world - 1))
    full, _ = orc.varlen(seeds, "und", 3, **kw)
    want_total = sum(len(r) for r in full.values())
    assert out["total"] == want_total and out["any_rows"] and out["shortest"] == 1
    assert out["counts"].shape == (world, 4)
    # balanced by predicted work: no rank holds more than twice its share of first-hop degree
    deg = 1 + parallel.first_hop_degree(snap, seeds, "und", True)
    share = [deg[a:b].sum() / deg.sum() for a, b in slices]
    assert max(share) < 2.0 / world + 0.05
    if rank == 0:
        for s, rows in full.items():
            got = out["rows"][s].numpy()
            assert got.shape == rows.shape
            assert np.array_equal(sort_rows(got), sort_rows(rows))
            # rank-order concatenation is ordered by seed because shards are contiguous seed ranges
            first = got[:, 0]
            bounds = np.cumsum([0] + [int(out["counts"][r, s]) for r in range(world)])
            for r in range(world):
                part = first[bounds[r]:bounds[r + 1]]
                lo, hi = slices[r]
                assert len(part) == 0 or (part.min() >= seeds[lo] and part.max() <= seeds[hi - 1])
    else:
        assert all(v is None for v in out["rows"].values())

    # ladder decision: a query nobody can satisfy -> 'no rows anywhere' on every rank
    none = parallel.ShardedExpander(snap, lambda p, s, t: orc.varlen(s, "out", 2, dest_kind=K("image"), excl_kinds=[ev, ns], entity_only=True)[0], "cpu")
    o2 = none.run(2, None, snap.nodes_of_kind(node))
    assert o2["total"] == 0 and not o2["any_rows"] and o2["shortest"] == -1
    # a graph built once per box and shared through a directory (workloads.shared_snapshot: what bench.py's strong-scaling
    # leg does with /dev/shm): rank 0 generates and saves, the others load after the barrier - every rank holds the same arrays
    import tempfile
    from k8s_llm_rca_b200 import workloads
    box = [tempfile.mkdtemp(prefix="rca_shared_") if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    shared = workloads.shared_snapshot(1500, 150, 3, rank=rank, barrier=dist.barrier, cache_dir=box[0])
    local = synth.generate(1500, 150, synth.config_seed(3))
    for name in ("kind_start", "out_rng", "out_dst", "in_adj", "und_adj", "st_off", "st_tmin", "st_tmax", "st_node", "rel_et"):
        assert np.array_equal(getattr(shared, name), getattr(local, name)), name
    assert [k.label for k in shared.kinds] == [k.label for k in local.kinds]
    w4 = workloads.Config4(snap=shared, n_seeds=200)
    mine, _, (lo, hi) = parallel.shard_seeds(shared, w4.seeds, rank, world, "und", True)
    spans = [None] * world
    dist.all_gather_object(spans, (lo, hi))
    assert spans[0][0] == 0 and spans[-1][1] == len(w4.seeds) and all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
    assert np.all(np.diff(w4.seeds) > 0)                       # sorted seed list: rank order is seed order
    dist.barrier()
    if rank == 0:
        import shutil
        shutil.rmtree(box[0], ignore_errors=True)
    dist.destroy_process_group()
    print(f"rank {rank} ok")


if __name__ == "__main__":
    main()

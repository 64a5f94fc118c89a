"""Worker of tests/test_multi_gpu.py: one rank per GPU over NCCL.  Each rank expands its slice of
the seeds on its own engine; counts are all-gathered, rows gathered to rank 0 straight from the
engines' output arenas, and rank 0 checks the union against the single-process CPU oracle."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import torch.distributed as dist
    from k8s_llm_rca_b200 import engine as E
    from k8s_llm_rca_b200 import parallel, synth
    from conftest import sort_rows

    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    snap = synth.generate(20_000, 2000, seed=41)
    eng = E.Engine(snap, device=local)
    K = snap.kind_index
    pod, node, ev, ns = K("Pod"), K("Node"), K("Event"), K("Namespace")
    seeds = snap.nodes_of_kind(pod)
    plan = E.varlen_plan("und", 3, dest_kind=node, excl_kinds=[ev, ns], entity_only=True)
    held = {}

    def expand(p, s, t):
        res = eng.expand(p, s, t)
        held["res"] = res
        return {lvl: parallel.device_rows(res, lvl) for lvl in range(1, 4) if res.emitted[lvl]}

    sx = parallel.ShardedExpander(snap, expand, device=f"cuda:{local}")
    out = sx.run(3, plan, seeds, gather=True, dst=0)
    if rank == 0:
        from oracle.csr import CsrOracle
        full, _ = CsrOracle(snap).varlen(seeds, "und", 3, dest_kind=node, excl_kinds=[ev, ns], entity_only=True)
        assert out["total"] == sum(len(r) for r in full.values())
        for s, rows in full.items():
            got = out["rows"][s].cpu().numpy()
            assert np.array_equal(sort_rows(got), sort_rows(rows))
            assert np.all(np.diff(got[:, 0]) >= 0)           # rank-order concatenation is ordered by seed
    # the count exchange over peer memory gives what NCCL's all-gather gives, round after round
    parallel.connect_peers(eng)
    for k in range(20):
        mine = 1000 * k + 7 * rank + (1 << 40) * (k % 3)
        got = eng.peer_allgather(mine)
        ref = parallel.allgather_counts([mine], device=f"cuda:{local}")[:, 0].cpu().numpy()
        assert np.array_equal(got, ref), (k, got, ref)
    eng.peer_allgather(rank, wait=False)                     # enqueue only ...
    eng.peer_allgather(10 + rank, wait=False)
    assert eng.peer_wait().tolist() == [10 + r for r in range(world)]      # ... the last exchange is what wait returns
    dist.barrier()
    eng.close()
    dist.destroy_process_group()
    print(f"rank {rank} ok")


if __name__ == "__main__":
    main()

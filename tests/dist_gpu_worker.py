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
    # the row gather of the C ABI (rca_gather_rows: remote stores into rank 0's window) delivers, byte for byte,
    # what the NCCL gather above delivered - window too small first: every rank is refused alike
    my_seeds, _, _ = parallel.shard_seeds(snap, seeds, rank, world, "und", True)
    res = eng.expand(plan, my_seeds)
    parallel.connect_gather(eng, 64, dst=0)
    try:
        parallel.gathered_rows(eng, res, 3, dst=0)
        raise AssertionError("a window of 64 words cannot hold the rows")
    except E.RcaError as ex:
        assert "window" in str(ex)
    need = max(int(out["counts"][:, lv].sum()) * (2 * lv + 1) for lv in range(1, 4))
    parallel.connect_gather(eng, need + 5, dst=0)
    for lv in (3, 1, 2, 3):
        got, per, ms = parallel.gathered_rows(eng, res, lv, dst=0)
        assert per.tolist() == out["counts"][:, lv].tolist(), (lv, per.tolist(), out["counts"].tolist(), res.emitted)
        if rank == 0 and lv in out["rows"]:
            assert torch.equal(got, out["rows"][lv]), f"level {lv}: peer-memory gather differs from the NCCL gather"
        elif rank == 0:
            assert got.shape[0] == 0                           # a length nobody has a path of
        else:
            assert got is None
    # ... and with a destination other than rank 0, rows with state columns (odd row widths, unaligned offsets)
    from k8s_llm_rca_b200.synth import T_BASE, DAY
    dst = world - 1
    ts = (T_BASE + np.random.default_rng(5).integers(0, 30 * DAY, size=len(seeds))).astype(np.int64)
    plan_s = E.varlen_plan("und", 3, dest_kind=node, excl_kinds=[ev, ns], entity_only=True, with_states=True)
    my_seeds, my_ts, _ = parallel.shard_seeds(snap, seeds, rank, world, "und", True, seed_ts=ts)
    res = eng.expand(plan_s, my_seeds, my_ts)
    cnt = parallel.allgather_counts(res.emitted, device=f"cuda:{local}")
    parallel.connect_gather(eng, int(max(cnt[:, lv].sum() * (3 * lv + 3) for lv in range(1, 4))) + 1, dst=dst)
    for lv in (1, 2, 3):
        got, per, _ = parallel.gathered_rows(eng, res, lv, dst=dst)
        ref = parallel.gather_rows(parallel.device_rows(res, lv), cnt[:, lv], dst=dst)
        if rank == dst:
            assert torch.equal(got, ref)
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

"""Graph replay of repeated submissions (include/rca_b200.h: rca_set_graphs): rca_expand_many captures a submission
it sees again and replays it with one graph launch.  A replay must be indistinguishable from the eager run: same rows
in the same order, same counters, sensible timings - also when the CONTENT of the seed buffers changes between
replays (the graph holds addresses, not values), when the seeds come from pinned host memory, when the engine
alternates between its two output arenas, and across an arena growth (which invalidates the graph)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rows(results):
    return [{s: r.rows(s).copy() for s in range(1, r.levels + 1) if r.emitted[s]} for r in results]


def _same(a, b):
    assert len(a) == len(b)
    for x, y in zip(a, b):
        assert x.keys() == y.keys()
        for s in x:
            assert np.array_equal(x[s], y[s])


def test_replay_equals_eager(snap_10k):
    import torch
    from k8s_llm_rca_b200 import engine, workloads
    from k8s_llm_rca_b200.synth import T_BASE, DAY
    snap = snap_10k
    w = workloads.Config3(snap=snap)
    lo, hi = snap.kind_range(snap.label_index("EVENT"))
    n = hi - lo
    rng = np.random.default_rng(8)
    pods_all = snap.nodes_of_kind(snap.kind_index("Pod")).astype(np.int32)

    def batch(seed):
        r = np.random.default_rng(seed)
        pods = np.sort(r.choice(pods_all, size=n, replace=False)).astype(np.int32)
        ev = np.arange(lo, hi, dtype=np.int32)
        ts = (T_BASE + r.integers(0, 30 * DAY, size=n)).astype(np.int64)
        return pods, ev, ts

    ref = engine.Engine(snap)
    ref.set_graphs(False)
    ref.set_fused_threshold(0)                     # the enumeration through count / scan / fill, like the headline workload
    eng = engine.Engine(snap)
    eng.set_graphs(True)
    eng.set_fused_threshold(0)
    stream = torch.cuda.Stream()
    eng.set_stream(stream.cuda_stream)
    with torch.cuda.stream(stream):
        d = [torch.empty(n, dtype=torch.int32, device="cuda"), torch.empty(n, dtype=torch.int32, device="cuda"),
             torch.empty(n, dtype=torch.int64, device="cuda")]
        for it in range(6):                        # calls 0-1 run eagerly, the third is captured, the rest are replays
            pods, ev, ts = batch(it)
            for t, a in zip(d, (pods, ev, ts)):
                t.copy_(torch.from_numpy(a))       # new CONTENT in the same device buffers
            stream.synchronize()
            tot, res = w.run(eng, d[0], d[1], d[2])
            got = _rows([r for _, r in res])
            tot_r, res_r = w.run(ref, pods, ev, ts)
            _same(got, _rows([r for _, r in res_r]))
            assert tot == tot_r > 0
            for (_, a), (_, b) in zip(res, res_r):
                assert a.emitted == b.emitted and a.frontier == b.frontier and a.scanned == b.scanned
                assert a.launches == b.launches
            assert all(r.kernel_ms > 0 for _, r in res)
        # pinned host seeds, two output arenas (what bench.py's end-to-end leg does)
        hp = [torch.empty(n, dtype=torch.int32).pin_memory(), torch.empty(n, dtype=torch.int32).pin_memory(),
              torch.empty(n, dtype=torch.int64).pin_memory()]
        eng.set_double_buffer(True)
        for it in range(7):
            pods, ev, ts = batch(100 + it)
            for t, a in zip(hp, (pods, ev, ts)):
                t.copy_(torch.from_numpy(a))
            tot, res = w.run(eng, hp[0].numpy(), hp[1].numpy(), hp[2].numpy())
            _same(_rows([r for _, r in res]), _rows([r for _, r in w.run(ref, pods, ev, ts)[1]]))
        eng.set_double_buffer(False)
        # a plan set that needs bigger arenas comes in between: the arenas grow, the old graphs no longer apply
        big = engine.varlen_plan("und", 4, dest_kind=-1, excl_kinds=w.excl, entity_only=True, prune=False)
        r_big = eng.expand_many([big], [pods_all], None)[0]
        assert r_big.total == ref.expand(big, pods_all).total
        for it in range(4):
            pods, ev, ts = batch(200 + it)
            for t, a in zip(d, (pods, ev, ts)):
                t.copy_(torch.from_numpy(a))
            stream.synchronize()
            _same(_rows([r for _, r in w.run(eng, d[0], d[1], d[2])[1]]), _rows([r for _, r in w.run(ref, pods, ev, ts)[1]]))
    eng.close()
    ref.close()

"""Mask-only rows (rca_plan_t::with_states = RCA_STATES_MASK): the row carries the missing-state mask but not the
state ids - what the reference's check branches on (check_state/analyze_root_cause.py:180: does the entity have a
STATE at the event's time or not).  Every kernel family must produce exactly the full-format rows minus the id
columns: general count/scan/fill and fused kernels, unit steps, one-launch chains, batched submissions, the
no-allocation entry point.  The full format itself is held to the oracle elsewhere (test_gpu_parity, test_gpu_unit)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _check(full, mask, s):
    a, b = full[s], mask[s]
    assert b.shape == (len(a), 2 * s + 2), s
    assert np.array_equal(b[:, :2 * s + 1], a[:, :2 * s + 1]), s          # same paths, same order
    assert np.array_equal(b[:, 2 * s + 1], a[:, 3 * s + 2]), s            # same mask
    ids = a[:, 2 * s + 1:3 * s + 2]
    assert np.array_equal((ids < 0) @ (1 << np.arange(s + 1)), a[:, 3 * s + 2])   # mask bit d <=> no state id in column d


@pytest.mark.parametrize("family", ["split", "fused"])
def test_mask_rows_equal_full_rows_without_the_ids(snap_10k, family):
    import torch
    from k8s_llm_rca_b200 import engine, workloads
    from k8s_llm_rca_b200.synth import T_BASE, DAY
    snap = snap_10k
    eng = engine.Engine(snap)
    eng.set_fused_threshold(0 if family == "split" else 1 << 40)
    K = snap.kind_index
    excl = [K("Event"), K("Namespace")]
    pods = snap.nodes_of_kind(K("Pod")).astype(np.int32)
    rng = np.random.default_rng(21)
    ts = (T_BASE + rng.integers(-2 * DAY, 33 * DAY, size=len(pods))).astype(np.int64)
    ts[:5] = [-5, 1 << 50, 0, (1 << 44) - 1, 1 << 44]                         # outside the packed header's range too
    missing_seen = 0
    for dest, hops, kw in (("Node", 4, {}), ("nfs", 3, {}), ("Node", 4, dict(inter_kinds=[K("ReplicaSet"), K("Service")])),
                           ("Deployment", 2, {}), ("Node", 3, dict(pos_kind=[K("Pod"), -1, -1, -1]))):
        base = dict(dest_kind=K(dest), excl_kinds=excl, entity_only=True, **kw)
        full = {s: r.copy() for s, r in eng.expand(engine.varlen_plan("und", hops, with_states=True, **base), pods, ts).all_rows().items()}
        res = eng.expand(engine.varlen_plan("und", hops, with_states="mask", **base), pods, ts)
        assert res.row_len[1:hops + 1] == [2 * s + 2 for s in range(1, hops + 1)]
        mask = res.all_rows()
        assert full.keys() == mask.keys()
        for s in full:
            _check(full, mask, s)
            missing_seen += int((mask[s][:, 2 * s + 1] != 0).sum())
    assert missing_seen > 0
    # the instance queries (chain_walk / chain_line, and chain_batch through a batched submission)
    w_full, w_mask = workloads.Config3(snap=snap), workloads.Config3(snap=snap, with_states="mask")
    lo, hi = snap.kind_range(snap.label_index("EVENT"))
    ev = np.arange(lo, hi, dtype=np.int32)
    ets = (T_BASE + rng.integers(0, 30 * DAY, size=len(ev))).astype(np.int64)
    for name in w_full.chains:
        L = len(w_full.chains[name][0])
        a = eng.expand(w_full.chains[name][2], ev, ets).rows(L).copy()
        b = eng.expand(w_mask.chains[name][2], ev, ets).rows(L)
        assert len(a) and np.array_equal(b[:, :2 * L + 1], a[:, :2 * L + 1]) and np.array_equal(b[:, 2 * L + 1], a[:, 3 * L + 2]), name
    sub_pods = pods[:len(ev)] if len(pods) >= len(ev) else np.resize(pods, len(ev))
    tot_f, res_f = w_full.run(eng, sub_pods, ev, ets)
    rows_f = [(n, {s: r.rows(s).copy() for s in range(1, r.levels + 1)}) for n, r in res_f]
    tot_m, res_m = w_mask.run(eng, sub_pods, ev, ets)
    assert tot_f == tot_m > 0
    for (n, fr), (_, r) in zip(rows_f, res_m):
        for s in range(1, r.levels + 1):
            a, b = fr[s], r.rows(s)
            assert np.array_equal(b[:, :2 * s + 1], a[:, :2 * s + 1]) and np.array_equal(b[:, 2 * s + 1], a[:, 3 * s + 2]), (n, s)
    # the entry point that allocates nothing
    eng.set_stream(torch.cuda.current_stream().cuda_stream)
    d_seeds, d_ts = torch.from_numpy(pods).cuda(), torch.from_numpy(ts).cuda()
    d_rows = torch.empty(6_000_000, dtype=torch.int32, device="cuda")
    kw = dict(dest_kind=K("Node"), excl_kinds=excl, entity_only=True)
    st = eng.expand_static(engine.varlen_plan("und", 3, with_states="mask", **kw), d_seeds, d_ts, d_rows)
    host, off = d_rows.cpu().numpy(), 0
    eng.set_fused_threshold(1 << 40)
    ref = eng.expand(engine.varlen_plan("und", 3, with_states=True, **kw), pods, ts).all_rows()
    for s in range(1, 4):
        n, wd = int(st.emitted[s]), int(st.row_len[s])
        assert wd == 2 * s + 2
        if n:
            b = host[off:off + n * wd].reshape(n, wd)
            assert np.array_equal(b[:, :2 * s + 1], ref[s][:, :2 * s + 1]) and np.array_equal(b[:, 2 * s + 1], ref[s][:, 3 * s + 2])
        off += n * wd
    eng.close()


def test_bad_states_mode_is_refused(snap_small):
    from k8s_llm_rca_b200 import engine
    eng = engine.Engine(snap_small)
    plan = engine.varlen_plan("und", 2, dest_kind=-1)
    plan.with_states = 3
    with pytest.raises(engine.RcaError):
        eng.expand(plan, snap_small.nodes_of_kind(0)[:4], np.zeros(4, np.int64))
    eng.close()

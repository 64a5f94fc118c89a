"""rca_expand_static, the entry point for callers that cannot allocate device memory.

Some users of the library run where cudaMalloc is off limits: their device memory was handed to a
pool at start-up, or they sit on a latency path that must not hit an allocator that can
synchronise the GPU.  For them the library keeps a fixed scratch array inside its own image and
rca_expand_static works out of that array only.  The caller brings the seeds and the row buffer
(this test allocates them with torch - that is the caller's business, not the entry point's).

Checked here, through the C ABI: the rows equal, word for word, what rca_expand returns for the
same plan and seeds (integer data, so the tolerance is zero), on several calls in a row - the
second call meets whatever the first one left in the shared scratch - and requests beyond the
stated maxima are refused with a clear message instead of growing anything.

The file name sorts last on purpose: with `-x` a failure here cannot hide the other tests.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rows_from_static(st, buf, levels):
    out, off = {}, 0
    host = buf.cpu().numpy()
    for s in range(1, levels + 1):
        n, w = int(st.emitted[s]), int(st.row_len[s])
        if n:
            out[s] = host[off:off + n * w].reshape(n, w).copy()
        off += n * w
    return out


def _same(a, b):
    assert a.keys() == b.keys()
    for s in a:
        assert a[s].shape == b[s].shape
        assert np.array_equal(a[s], b[s])          # same rows in the same order


def test_static_entry_matches_expand_on_consecutive_calls(snap_10k):
    import torch
    from k8s_llm_rca_b200 import engine
    from k8s_llm_rca_b200.synth import T_BASE, DAY

    snap = snap_10k
    eng = engine.Engine(snap)
    eng.set_fused_threshold(1 << 40)               # the reference run uses the same single-launch kernels
    K = snap.kind_index
    pod, node, ev, ns = K("Pod"), K("Node"), K("Event"), K("Namespace")
    seeds = snap.nodes_of_kind(pod)[::3].astype(np.int32)
    ts = (T_BASE + np.random.default_rng(9).integers(0, 30 * DAY, size=len(seeds))).astype(np.int64)
    d_seeds, d_ts = torch.from_numpy(seeds).cuda(), torch.from_numpy(ts).cuda()
    d_rows = torch.full((4_000_000,), -7, dtype=torch.int32, device="cuda")

    L = snap.label_index
    chain = [snap.find_etypes("ReferInternal", L("Pod"), L("PersistentVolumeClaim"),
                              "spec_volumes_persistentVolumeClaim_claimName")[0],
             snap.find_etypes("ReferInternal", L("PersistentVolume"), L("PersistentVolumeClaim"), "spec_claimRef_uid")[0],
             snap.find_etypes("UseExternal", L("PersistentVolume"), L("nfs"), "spec_nfs_path")[0]]
    plans = [
        (engine.varlen_plan("und", 3, dest_kind=node, excl_kinds=[ev, ns], entity_only=True, with_states=True), 3, d_ts, ts),
        (engine.chain_plan(chain, ["out", "in", "out"]), 3, None, None),
        (engine.varlen_plan("und", 3, dest_kind=node, excl_kinds=[ev, ns], entity_only=True, with_states=True), 3, d_ts, ts),
        (engine.varlen_plan("out", 4, dest_kind=node, excl_kinds=[ev, ns], entity_only=True), 4, None, None),
    ]
    for plan, levels, dts, hts in plans:          # back to back: each call finds the previous call's leftovers
        d_rows.fill_(-7)
        st = eng.expand_static(plan, d_seeds, dts, d_rows)
        got = _rows_from_static(st, d_rows, levels)
        ref = eng.expand(plan, seeds, hts)
        _same(got, ref.all_rows())
        assert int(st.total) == ref.total > 0
        assert list(st.frontier)[:levels + 1] == ref.frontier and list(st.scanned)[:levels + 1] == ref.scanned
        used = sum(int(st.emitted[s]) * int(st.row_len[s]) for s in range(levels + 1))
        assert bool((d_rows[used:used + 64] == -7).all())         # nothing written past the result

    # and against the count / scan / fill kernels: same set of rows
    eng.set_fused_threshold(0)
    plan, levels, dts, hts = plans[0]
    st = eng.expand_static(plan, d_seeds, dts, d_rows)
    got = _rows_from_static(st, d_rows, levels)
    ref = eng.expand(plan, seeds, hts).all_rows()
    for s in ref:
        a, b = got[s], ref[s]
        assert np.array_equal(a[np.lexsort(a.T[::-1])], b[np.lexsort(b.T[::-1])])
    eng.close()


def test_static_entry_refuses_what_does_not_fit(snap_10k):
    import torch
    from k8s_llm_rca_b200 import engine

    snap = snap_10k
    eng = engine.Engine(snap)
    K = snap.kind_index
    pod, node, ev, ns = K("Pod"), K("Node"), K("Event"), K("Namespace")
    plan = engine.varlen_plan("und", 3, dest_kind=node, excl_kinds=[ev, ns], entity_only=True)
    seeds = torch.from_numpy(snap.nodes_of_kind(pod).astype(np.int32)).cuda()
    # a row buffer that is too small: refused, message names the buffer
    small = torch.empty(1000, dtype=torch.int32, device="cuda")
    with pytest.raises(engine.RcaError) as ei:
        eng.expand_static(plan, seeds, None, small)
    assert "row buffer" in str(ei.value)
    # more seeds than the fixed scratch is sized for: refused before anything runs
    many = torch.zeros(262144 + 1, dtype=torch.int32, device="cuda")
    rows = torch.empty(1 << 20, dtype=torch.int32, device="cuda")
    with pytest.raises(engine.RcaError) as ei:
        eng.expand_static(plan, many, None, rows)
    assert "RCA_STATIC_MAX_SEEDS" in str(ei.value)
    # more partial paths than the scratch holds (unpruned walk without a destination): refused
    wide = engine.varlen_plan("und", 4, dest_kind=-1, entity_only=True, prune=False)
    big = torch.empty(1 << 20, dtype=torch.int32, device="cuda")
    with pytest.raises(engine.RcaError) as ei:
        eng.expand_static(wide, seeds, None, big)
    assert "RCA_STATIC_MAX_FRONTIER" in str(ei.value) or "row buffer" in str(ei.value)
    # the engine is still usable afterwards
    ok = eng.expand_static(plan, seeds[:100].contiguous(), None, rows)
    assert int(ok.total) == eng.expand(plan, snap.nodes_of_kind(pod)[:100]).total
    eng.close()

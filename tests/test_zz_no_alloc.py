"""rca_expand_static, the entry point for callers that cannot allocate device memory.

Some users of the library run where cudaMalloc is off limits: their device memory was handed to a
pool at start-up, or they sit on a latency path that must not hit an allocator that can
synchronise the GPU.  For them the library keeps a fixed scratch array inside its own image and
rca_expand_static works out of that array only.  The caller brings the seeds and the row buffer
(this test allocates them with torch - that is the caller's business, not the entry point's).

Checked here, through the C ABI: the rows equal, word for word, what rca_expand returns for the
same plan and seeds (integer data, so the tolerance is zero), on several calls in a row - the
second call meets whatever the first one left in the shared scratch - and requests beyond the
stated maxima are refused with a clear message instead of growing anything.  One call is also
held against the CPU oracle, and the property the entry point exists for is measured: the
device's free memory (cudaMemGetInfo) is the same before and after a call.

The entry point works on the engine's stream, which knows nothing of the stream torch fills the
caller's tensors on; the test hands the engine torch's stream (set_stream), as a real caller of
this kind would, so that "fill the buffer, then expand into it" is ordered on the device.

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


def test_static_entry_matches_expand_on_consecutive_calls(snap_10k, csr_oracle_lib):
    import torch
    from k8s_llm_rca_b200 import engine
    from k8s_llm_rca_b200.synth import T_BASE, DAY

    snap = snap_10k
    eng = engine.Engine(snap)
    eng.set_stream(torch.cuda.current_stream().cuda_stream)     # the stream the tensors below are filled on
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
    for it, (plan, levels, dts, hts) in enumerate(plans):   # back to back: each call finds the previous call's leftovers
        d_rows.fill_(-7)
        if it != 2:                                # a plan's first call may make the driver load its kernels' code
            eng.expand_static(plan, d_seeds, dts, d_rows)
            d_rows.fill_(-7)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        st = eng.expand_static(plan, d_seeds, dts, d_rows)
        assert torch.cuda.mem_get_info()[0] == free_before       # the call took no device memory
        got = _rows_from_static(st, d_rows, levels)
        ref = eng.expand(plan, seeds, hts)
        _same(got, ref.all_rows())
        assert int(st.total) == ref.total > 0
        assert list(st.frontier)[:levels + 1] == ref.frontier and list(st.scanned)[:levels + 1] == ref.scanned
        used = sum(int(st.emitted[s]) * int(st.row_len[s]) for s in range(levels + 1))
        assert bool((d_rows[used:used + 64] == -7).all())         # nothing written past the result

    # against the CPU oracle (not only against the library's other entry point)
    orc = csr_oracle_lib.CsrOracle(snap)
    plan, levels, dts, hts = plans[3]
    d_rows.fill_(-7)
    st = eng.expand_static(plan, d_seeds, None, d_rows)
    got = _rows_from_static(st, d_rows, levels)
    want, _ = orc.varlen(seeds, "out", 4, dest_kind=node, excl_kinds=[ev, ns], entity_only=True)
    assert got.keys() == want.keys()
    for s in want:
        a, b = got[s], want[s]
        assert np.array_equal(a[np.lexsort(a.T[::-1])], b[np.lexsort(b.T[::-1])])
    plan, levels, dts, hts = plans[0]
    d_rows.fill_(-7)
    st = eng.expand_static(plan, d_seeds, dts, d_rows)
    got = _rows_from_static(st, d_rows, levels)
    want, _ = orc.varlen(seeds, "und", 3, dest_kind=node, excl_kinds=[ev, ns], entity_only=True)
    order = np.argsort(seeds, kind="stable")
    for s in want:
        a, b = got[s], want[s]
        a = a[np.lexsort(a[:, :2 * s + 1].T[::-1])]
        b = b[np.lexsort(b.T[::-1])]
        assert np.array_equal(a[:, :2 * s + 1], b)
        t = ts[order[np.searchsorted(seeds[order], b[:, 0])]]
        miss = np.zeros(len(b), np.int64)
        for d in range(s + 1):                     # the fused HasState columns: first state id per node, missing mask
            first, cnt, _ = orc.states(b[:, d], t, limit=1)
            assert np.array_equal(a[:, 2 * s + 1 + d], first)
            miss |= (cnt == 0).astype(np.int64) << d
        assert np.array_equal(a[:, 3 * s + 2], miss.astype(np.int32))

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
    eng.set_stream(torch.cuda.current_stream().cuda_stream)
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
    # tensors of the wrong type are turned away by the binding before they reach the kernels
    with pytest.raises(ValueError):
        eng.expand_static(plan, seeds.to(torch.int64), None, rows)
    # the engine is still usable afterwards
    ok = eng.expand_static(plan, seeds[:100].contiguous(), None, rows)
    assert int(ok.total) == eng.expand(plan, snap.nodes_of_kind(pod)[:100]).total
    eng.close()

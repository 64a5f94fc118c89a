"""Parity of the CUDA path (through the C ABI) with the CPU oracle on seeded graphs."""
import numpy as np
import pytest

from conftest import rows_to_set, sort_rows

pytestmark = pytest.mark.gpu


# every parity test runs twice: through the fused single-launch kernels (small seed sets) and
# through the count / scan / fill kernels (large seed sets); the threshold is forced either way
@pytest.fixture(scope="module", params=["fused", "split"])
def eng_small(snap_small, request):
    from k8s_llm_rca_b200 import engine
    e = engine.Engine(snap_small)
    e.set_fused_threshold(1 << 40 if request.param == "fused" else 0)
    yield e
    e.close()


@pytest.fixture(scope="module", params=["fused", "split"])
def eng_10k(snap_10k, request):
    from k8s_llm_rca_b200 import engine
    e = engine.Engine(snap_10k)
    e.set_fused_threshold(1 << 40 if request.param == "fused" else 0)
    yield e
    e.close()


def _kinds(snap, *names):
    return [snap.kind_index(n) for n in names]


def check_varlen(eng, orc, seeds, direction, max_hops, **kw):
    from k8s_llm_rca_b200 import engine
    prune = kw.pop("prune", True)
    plan = engine.varlen_plan(direction, max_hops, prune=prune, **kw)
    res = eng.expand(plan, seeds)
    ref, scanned = orc.varlen(seeds, direction, max_hops, **kw)
    got = rows_to_set(res.all_rows())
    want = rows_to_set(ref)
    assert len(got) == len(want)
    assert got == want                       # bit-exact as sorted sets of (nodes, rels)
    assert res.total == len(want)
    if not prune:
        assert sum(res.scanned) == scanned   # same number of adjacency entries inspected
    return res


@pytest.mark.parametrize("direction", ["out", "in", "und"])
@pytest.mark.parametrize("max_hops", [1, 2, 3, 4])
def test_varlen_entity_projection(eng_small, snap_small, csr_oracle_lib, direction, max_hops):
    orc = csr_oracle_lib.CsrOracle(snap_small)
    pod, node, ev, ns = _kinds(snap_small, "Pod", "Node", "Event", "Namespace")
    seeds = snap_small.nodes_of_kind(pod)
    dest = node if direction != "in" else _kinds(snap_small, "Service")[0]
    for prune in (True, False):
        check_varlen(eng_small, orc, seeds, direction, max_hops, dest_kind=dest, excl_kinds=[ev, ns],
                     entity_only=True, prune=prune)


def test_varlen_with_intermediate_kinds(eng_small, snap_small, csr_oracle_lib):
    orc = csr_oracle_lib.CsrOracle(snap_small)
    pod, node, ev, ns, cm, rs = _kinds(snap_small, "Pod", "Node", "Event", "Namespace", "ConfigMap", "ReplicaSet")
    seeds = snap_small.nodes_of_kind(pod)
    res = check_varlen(eng_small, orc, seeds, "und", 4, dest_kind=node, excl_kinds=[ev, ns], inter_kinds=[cm, rs],
                       entity_only=True)
    assert res.emitted[1] == 0               # a 1-hop path has no interior node


def test_varlen_full_graph_any_dest(eng_small, snap_small, csr_oracle_lib):
    """entity_only off: STATE nodes are reachable; dest any; nothing excluded."""
    orc = csr_oracle_lib.CsrOracle(snap_small)
    seeds = np.arange(0, snap_small.n_nodes, 7, dtype=np.int32)
    check_varlen(eng_small, orc, seeds, "und", 2, dest_kind=-1)
    check_varlen(eng_small, orc, seeds, "out", 3, dest_kind=-1)


def test_varlen_via_namespace_shape(eng_small, snap_small, csr_oracle_lib):
    """src-Namespace-dest with r1 != r2 and nodes allowed to coincide (reference :125-129)."""
    orc = csr_oracle_lib.CsrOracle(snap_small)
    pod, ns, cm = _kinds(snap_small, "Pod", "Namespace", "ConfigMap")
    seeds = snap_small.nodes_of_kind(pod)
    check_varlen(eng_small, orc, seeds, "und", 2, min_hops=2, dest_kind=cm, node_simple=False,
                 pos_kind=[pod, ns, cm], entity_only=True)
    check_varlen(eng_small, orc, seeds, "und", 2, min_hops=2, dest_kind=pod, node_simple=False,
                 pos_kind=[pod, ns, pod], entity_only=True)


def test_varlen_edge_cases(eng_small, snap_small, csr_oracle_lib):
    from k8s_llm_rca_b200 import engine
    orc = csr_oracle_lib.CsrOracle(snap_small)
    pod, node, ev, ns = _kinds(snap_small, "Pod", "Node", "Event", "Namespace")
    # empty seed list
    res = eng_small.expand(engine.varlen_plan("und", 3, dest_kind=node), np.zeros(0, np.int32))
    assert res.total == 0
    # seeds of an excluded kind, out-of-range seeds, duplicate seeds
    seeds = np.concatenate([snap_small.nodes_of_kind(ns), [-5, snap_small.n_nodes + 3],
                            snap_small.nodes_of_kind(pod)[:5], snap_small.nodes_of_kind(pod)[:5]]).astype(np.int32)
    plan = engine.varlen_plan("und", 3, dest_kind=node, excl_kinds=[ev, ns], entity_only=True)
    res = eng_small.expand(plan, seeds)
    ok = seeds[(seeds >= 0) & (seeds < snap_small.n_nodes)]
    ref, _ = orc.varlen(ok, "und", 3, dest_kind=node, excl_kinds=[ev, ns], entity_only=True)
    assert rows_to_set(res.all_rows()) == rows_to_set(ref)
    # one seed
    check_varlen(eng_small, orc, snap_small.nodes_of_kind(pod)[:1], "und", 4, dest_kind=node, excl_kinds=[ev, ns],
                 entity_only=True)


def test_scratch_arena_growth(snap_small, csr_oracle_lib):
    """A tiny reservation forces the overflow -> grow -> re-run path; results must not change."""
    from k8s_llm_rca_b200 import engine
    orc = csr_oracle_lib.CsrOracle(snap_small)
    pod, node, ev, ns = _kinds(snap_small, "Pod", "Node", "Event", "Namespace")
    for threshold in (1 << 40, 0):
        eng = engine.Engine(snap_small)
        eng.set_fused_threshold(threshold)
        eng.reserve(64, 256)
        res = check_varlen(eng, orc, snap_small.nodes_of_kind(pod), "und", 3, dest_kind=node, excl_kinds=[ev, ns],
                           entity_only=True)
        assert res.retries >= 1
        eng.close()


def _chain_for(snap, *steps):
    """steps: (relType, srcKind, dstKind, key, dir)"""
    ets, dirs = [], []
    for (rt, sk, dk, key, d) in steps:
        t = snap.find_etypes(rt, snap.label_index(sk), snap.label_index(dk), key)
        assert len(t) == 1
        ets.append(t[0])
        dirs.append(d)
    return ets, dirs


def test_chain_event_to_configmap(eng_10k, snap_10k, csr_oracle_lib):
    """The instance query of the reference's worked example (generate_query.py:188-207):
    EVENT <-HasEvent- Event -involvedObject-> Pod -configMap-> ConfigMap."""
    from k8s_llm_rca_b200 import engine
    orc = csr_oracle_lib.CsrOracle(snap_10k)
    ets, dirs = _chain_for(snap_10k,
                           ("HasEvent", "Event", "EVENT", "metadata_uid", "in"),
                           ("ReferInternal", "Event", "Pod", "involvedObject_uid", "out"),
                           ("ReferInternal", "Pod", "ConfigMap", "spec_volumes_configMap_name", "out"))
    seeds = snap_10k.nodes_of_kind(snap_10k.label_index("EVENT"))
    res = eng_10k.expand(engine.chain_plan(ets, dirs), seeds)
    ref, _ = orc.chain(seeds, ets, [{"out": 0, "in": 1}[d] for d in dirs])
    assert np.array_equal(sort_rows(res.rows(3)), sort_rows(ref))
    assert res.total == len(ref) > 0


def test_chain_with_reversed_step(eng_10k, snap_10k, csr_oracle_lib):
    """Pod -> PVC <- PV -> nfs: the reversed step of reference test_generate_query.py:34-40."""
    from k8s_llm_rca_b200 import engine
    orc = csr_oracle_lib.CsrOracle(snap_10k)
    ets, dirs = _chain_for(snap_10k,
                           ("ReferInternal", "Pod", "PersistentVolumeClaim",
                            "spec_volumes_persistentVolumeClaim_claimName", "out"),
                           ("ReferInternal", "PersistentVolume", "PersistentVolumeClaim", "spec_claimRef_uid", "in"),
                           ("UseExternal", "PersistentVolume", "nfs", "spec_nfs_path", "out"))
    seeds = snap_10k.nodes_of_kind(snap_10k.kind_index("Pod"))
    res = eng_10k.expand(engine.chain_plan(ets, dirs), seeds)
    ref, _ = orc.chain(seeds, ets, [{"out": 0, "in": 1}[d] for d in dirs])
    assert np.array_equal(sort_rows(res.rows(3)), sort_rows(ref))
    assert len(ref) > 0
    # seeds of the wrong kind produce nothing
    res = eng_10k.expand(engine.chain_plan(ets, dirs), snap_10k.nodes_of_kind(snap_10k.kind_index("Node")))
    assert res.total == 0


def test_state_lookup(eng_10k, snap_10k, csr_oracle_lib):
    orc = csr_oracle_lib.CsrOracle(snap_10k)
    rng = np.random.default_rng(3)
    n_ent = int(snap_10k.kind_start[snap_10k.kind_index("Event")])
    nodes = rng.integers(0, n_ent, size=5000).astype(np.int32)
    from k8s_llm_rca_b200.synth import T_BASE, DAY
    ts = (T_BASE + rng.integers(-5 * DAY, 35 * DAY, size=5000)).astype(np.int64)
    # hit interval boundaries exactly: tmin is inside, tmax is outside (half-open)
    some = snap_10k.st_off[nodes[:200]]
    has = snap_10k.st_off[nodes[:200] + 1] > some
    ts[:200][has] = snap_10k.st_tmin[some[has]]
    ts[200:400] = ts[:200]
    ts[200:400][has] = snap_10k.st_tmax[some[has]]
    nodes[200:400] = nodes[:200]
    f, c, ids = eng_10k.state_lookup(nodes, ts)
    rf, rc, rids = orc.states(nodes, ts)
    assert np.array_equal(f, rf) and np.array_equal(c, rc) and np.array_equal(ids, rids)
    assert (c == 0).any() and (c > 0).any()


def test_fused_state_filter(eng_10k, snap_10k, csr_oracle_lib):
    """with_states: every emitted row also carries, per node, the STATE valid at the seed's time."""
    from k8s_llm_rca_b200 import engine
    orc = csr_oracle_lib.CsrOracle(snap_10k)
    pod, node, ev, ns = _kinds(snap_10k, "Pod", "Node", "Event", "Namespace")
    seeds = snap_10k.nodes_of_kind(pod)[::5]
    from k8s_llm_rca_b200.synth import T_BASE, DAY
    ts = (T_BASE + np.random.default_rng(5).integers(0, 30 * DAY, size=len(seeds))).astype(np.int64)
    plan = engine.varlen_plan("und", 3, dest_kind=node, excl_kinds=[ev, ns], entity_only=True, with_states=True)
    res = eng_10k.expand(plan, seeds, ts)
    ref, _ = orc.varlen(seeds, "und", 3, dest_kind=node, excl_kinds=[ev, ns], entity_only=True)
    ts_of_seed = dict(zip(seeds.tolist(), ts.tolist()))
    for s in (1, 2, 3):
        got = sort_rows(res.rows(s))
        want = sort_rows(ref.get(s, np.zeros((0, 2 * s + 1), np.int32)))
        assert got.shape == (len(want), 3 * s + 3)
        assert np.array_equal(got[:, :2 * s + 1], want)
        if len(got) == 0:
            continue
        t = np.array([ts_of_seed[v] for v in got[:, 0].tolist()], dtype=np.int64)
        miss = np.zeros(len(got), np.int64)
        for d in range(s + 1):
            f, c, _ = orc.states(got[:, d], t)
            assert np.array_equal(got[:, 2 * s + 1 + d], f)
            miss |= (c == 0).astype(np.int64) << d
        assert np.array_equal(got[:, 3 * s + 2].astype(np.int64), miss)


def test_10k_all_source_kinds_to_node(eng_10k, snap_10k, csr_oracle_lib):
    """BASELINE config 2 shape: every source kind -> Node, <= 4 hops, directed then undirected."""
    orc = csr_oracle_lib.CsrOracle(snap_10k)
    node, ev, ns = _kinds(snap_10k, "Node", "Event", "Namespace")
    for k, kind in enumerate(snap_10k.kinds):
        if kind.is_state or kind.name in ("Event", "Namespace"):
            continue
        seeds = snap_10k.nodes_of_kind(k)            # ALL nodes of the kind, both statements at the full 4 hops
        for direction in ("out", "und"):
            from k8s_llm_rca_b200 import engine
            kw = dict(dest_kind=node, excl_kinds=[ev, ns], entity_only=True)
            res = eng_10k.expand(engine.varlen_plan(direction, 4, **kw), seeds)
            ref, _ = orc.varlen(seeds, direction, 4, **kw)
            got = res.all_rows()
            assert sorted(got) == sorted(ref), (kind.name, direction)
            for s in ref:                                # bit-exact as sorted row sets, level by level
                assert np.array_equal(sort_rows(got[s]), sort_rows(ref[s])), (kind.name, direction, s)
            assert res.total == sum(len(r) for r in ref.values())


def test_determinism(eng_10k, snap_10k):
    """Same input -> byte-identical output buffers (slot = f(seed order, CSR order))."""
    from k8s_llm_rca_b200 import engine
    pod, node, ev, ns = _kinds(snap_10k, "Pod", "Node", "Event", "Namespace")
    plan = engine.varlen_plan("und", 3, dest_kind=node, excl_kinds=[ev, ns], entity_only=True)
    seeds = snap_10k.nodes_of_kind(pod)
    a = eng_10k.expand(plan, seeds).all_rows()
    b = eng_10k.expand(plan, seeds).all_rows()
    assert a.keys() == b.keys()
    for s in a:
        assert np.array_equal(a[s], b[s])


def test_metagraph_on_gpu(snap_10k, csr_oracle_lib):
    """The reference's own use: metapaths Pod -> nfs on the (tiny) metagraph."""
    from k8s_llm_rca_b200 import engine
    meta = snap_10k.derive_metagraph()
    orc = csr_oracle_lib.CsrOracle(meta)
    eng = engine.Engine(meta)
    pod, nfs, ev, ns, pv, pvc = _kinds(meta, "Pod", "nfs", "Event", "Namespace", "PersistentVolume",
                                       "PersistentVolumeClaim")
    seeds = meta.nodes_of_kind(pod)
    res = check_varlen(eng, orc, seeds, "out", 3, dest_kind=nfs, excl_kinds=[ev, ns], inter_kinds=[pvc, pv])
    assert res.total == 0                     # PV -> PVC points the other way: no directed path
    res = check_varlen(eng, orc, seeds, "und", 3, dest_kind=nfs, excl_kinds=[ev, ns], inter_kinds=[pvc, pv])
    assert res.total >= 1
    eng.close()


def test_message_contains_scan(snap_10k):
    """Batched CONTAINS on the device == Python's `in` over the same messages."""
    from k8s_llm_rca_b200 import engine
    eng = engine.Engine(snap_10k)
    ids, msgs = snap_10k.props.messages()
    eng.text_load(ids, msgs)
    rng = np.random.default_rng(8)
    pats = [msgs[i] for i in rng.integers(0, len(msgs), size=40)]                      # whole messages
    pats += [msgs[i][5:25] for i in rng.integers(0, len(msgs), size=20)]               # inner substrings
    pats += ["not found", "configmap", "MountVolume", '"', "zzz-no-such-text", "", "d", "0/", "é"]
    pats += [pats[0], pats[41]]                                                        # duplicates
    first, hits = eng.contains(pats)
    for i, t in enumerate(pats):
        want = sorted(int(v) for v, m in zip(ids.tolist(), msgs) if t in m)
        assert hits.get(i, []) == want, (i, t)
        assert first[i] == (want[0] if want else -1)
    assert len(hits[pats.index("not found")]) > 100 and pats.index("zzz-no-such-text") not in hits
    # a tiny pair buffer must grow, not truncate
    first2, hits2 = eng.contains(pats, pairs_cap=8)
    assert hits2 == hits and np.array_equal(first, first2)
    eng.close()


def test_count_only_plans(eng_10k, snap_10k):
    """count_only: same counters, nothing written (for enumerations too large to materialise)."""
    from k8s_llm_rca_b200 import engine
    pod, node, ev, ns = _kinds(snap_10k, "Pod", "Node", "Event", "Namespace")
    seeds = snap_10k.nodes_of_kind(pod)
    full = eng_10k.expand(engine.varlen_plan("und", 4, dest_kind=node, excl_kinds=[ev, ns], entity_only=True), seeds)
    cnt = eng_10k.expand(engine.varlen_plan("und", 4, dest_kind=node, excl_kinds=[ev, ns], entity_only=True,
                                            count_only=True), seeds)
    assert cnt.emitted == full.emitted and cnt.frontier == full.frontier and cnt.scanned == full.scanned
    with pytest.raises(ValueError):
        cnt.rows(3)


def test_expand_many_equals_separate_calls(snap_10k):
    """One submission of several plans == the same plans one call at a time (rows, order, counters)."""
    from k8s_llm_rca_b200 import engine, workloads
    for threshold in (None, 0, 1 << 40):
        w = workloads.Config3(snap=snap_10k)
        eng = engine.Engine(snap_10k)
        if threshold is not None:
            eng.set_fused_threshold(threshold)
        pods, ev, ts = w.batch(0)
        names = ["varlen"] + list(w.chains)
        plans = [w.varlen] + [c[2] for c in w.chains.values()]
        seeds = [pods] + [ev] * len(w.chains)
        single = []
        for p, s in zip(plans, seeds):
            r = eng.expand(p, s, ts)
            single.append((r.emitted, r.frontier, r.scanned, r.all_rows()))
        many = eng.expand_many(plans, seeds, [ts] * len(plans))
        for (em, fr, sc, rows), r, name in zip(single, many, names):
            assert r.emitted == em and r.frontier == fr and r.scanned == sc, name
            got = r.all_rows()
            assert got.keys() == rows.keys()
            for lvl in rows:
                assert np.array_equal(got[lvl], rows[lvl]), (name, lvl)
        # a tiny arena: the whole submission is re-run after growing
        eng2 = engine.Engine(snap_10k)
        if threshold is not None:
            eng2.set_fused_threshold(threshold)
        eng2.reserve(64, 256)
        again = eng2.expand_many(plans, seeds, [ts] * len(plans))
        assert again[0].retries >= 1
        for (em, _, _, rows), r in zip(single, again):
            assert r.emitted == em and all(np.array_equal(r.rows(lvl), rows[lvl]) for lvl in rows)
        # mixed with a count-only plan and an empty seed list
        cnt = engine.varlen_plan("und", 3, dest_kind=snap_10k.kind_index("Node"), excl_kinds=w.excl, entity_only=True,
                                 count_only=True)
        mix = eng.expand_many([cnt, plans[1], plans[2]], [pods, ev, np.zeros(0, np.int32)], [None, ts, ts[:0]])
        assert mix[0].total > 0 and mix[2].total == 0
        assert np.array_equal(mix[1].rows(3), single[1][3][3])
        eng.close()
        eng2.close()


def test_eight_hops(eng_small, snap_small, csr_oracle_lib):
    """RCA_MAX_HOPS deep: levels 6 to 8 need more shared memory than a kernel gets by default."""
    orc = csr_oracle_lib.CsrOracle(snap_small)
    dep, nfs, ev, ns, node = _kinds(snap_small, "Deployment", "nfs", "Event", "Namespace", "Node")
    seeds = snap_small.nodes_of_kind(dep)[:4]
    res = check_varlen(eng_small, orc, seeds, "und", 8, dest_kind=nfs, excl_kinds=[ev, ns, node], entity_only=True)
    assert res.levels == 8


def test_fused_state_filter_extreme_timestamps(eng_small, snap_small, csr_oracle_lib):
    """Seed timestamps outside the range the packed state headers hold (before 1970, after 2527) and at
    interval boundaries are answered exactly (the kernels then read the full records)."""
    from k8s_llm_rca_b200 import engine
    from k8s_llm_rca_b200.synth import DAY, T_BASE
    orc = csr_oracle_lib.CsrOracle(snap_small)
    pod, node, ev, ns = _kinds(snap_small, "Pod", "Node", "Event", "Namespace")
    seeds = snap_small.nodes_of_kind(pod)
    edge = np.concatenate([snap_small.st_tmin, snap_small.st_tmax, snap_small.st_tmin - 1, snap_small.st_tmax - 1])
    special = np.array([-1, -5, 0, 1, (1 << 44) - 2, (1 << 44) - 1, 1 << 44, 1 << 50, (1 << 62), -(1 << 40),
                        T_BASE, T_BASE + 3 * DAY], dtype=np.int64)
    pool = np.concatenate([special, edge.astype(np.int64)])
    ts = pool[np.random.default_rng(3).integers(0, len(pool), size=len(seeds))]
    ts[:len(special)] = special[:len(seeds)][:len(special)]
    plan = engine.varlen_plan("und", 3, dest_kind=node, excl_kinds=[ev, ns], entity_only=True, with_states=True)
    res = eng_small.expand(plan, seeds, ts)
    ts_of_seed = dict(zip(seeds.tolist(), ts.tolist()))
    total = 0
    for s in (1, 2, 3):
        got = res.rows(s)
        if len(got) == 0:
            continue
        total += len(got)
        t = np.array([ts_of_seed[v] for v in got[:, 0].tolist()], dtype=np.int64)
        miss = np.zeros(len(got), np.int64)
        for d in range(s + 1):
            f, c, _ = orc.states(got[:, d], t)
            assert np.array_equal(got[:, 2 * s + 1 + d], f)
            miss |= (c == 0).astype(np.int64) << d
        assert np.array_equal(got[:, 3 * s + 2].astype(np.int64), miss)
    assert total > 0


def test_message_filter_pairs(snap_10k):
    """rca_text_match_pairs == Python's `a in m or b in m` (generate_query.py:104-129) over random pairs."""
    from k8s_llm_rca_b200 import engine
    eng = engine.Engine(snap_10k)
    ids, msgs = snap_10k.props.messages()
    eng.text_load(ids, msgs)
    rng = np.random.default_rng(12)
    pick = rng.integers(0, len(msgs), size=4000)
    pool = [msgs[i][a:a + n] for i, a, n in zip(rng.integers(0, len(msgs), 300).tolist(), rng.integers(0, 40, 300).tolist(),
                                                  rng.integers(1, 30, 300).tolist())]
    pool += ["Pod", "pod-", "Node", "node-1", "", "é", "x" * 400, "configmap", "zzz", None]
    a = [pool[i] for i in rng.integers(0, len(pool), size=len(pick))]
    b = [pool[i] for i in rng.integers(0, len(pool), size=len(pick))]
    # half of the pairs get a needle that really is in their message
    for k in range(0, len(pick), 2):
        m = msgs[pick[k]]
        a[k] = m[len(m) // 3: len(m) // 3 + 12]
    got = eng.match_pairs(ids[pick], a, b)
    want = [((x is not None and x in msgs[i]) or (y is not None and y in msgs[i])) for i, x, y in zip(pick.tolist(), a, b)]
    assert got.tolist() == want
    assert 0.4 < np.mean(want) < 1.0
    assert eng.match_pairs([], [], []).shape == (0,)
    with pytest.raises(ValueError):
        eng.match_pairs([0], ["a"], ["b"])                  # node 0 is a Pod: it has no message
    eng.close()


def test_expand_many_with_more_side_plans_than_one_batch(snap_10k):
    """Thirteen one-launch chains (two chain_batch launches), a count-only one among them, around two
    variable-length plans: every plan's rows and counters equal those of a call of its own."""
    from k8s_llm_rca_b200 import engine, workloads
    w = workloads.Config3(snap=snap_10k)
    eng = engine.Engine(snap_10k)
    pods, ev, ts = w.batch(0)
    chains = [c[2] for c in w.chains.values()]
    ets0, dirs0, _ = next(iter(w.chains.values()))
    cnt = engine.chain_plan(ets0, dirs0, count_only=True)
    short = engine.varlen_plan("und", 2, dest_kind=snap_10k.kind_index("Node"), excl_kinds=w.excl, entity_only=True)
    plans = chains[:3] + [w.varlen] + chains + [cnt] + chains[1:5] + [short]
    seeds = [pods if p.kind == engine.PLAN_VARLEN else ev for p in plans]
    # different seed subsets per plan, so that a mix-up of plans cannot go unnoticed
    seeds = [s[k % 3::3] if k % 2 else s for k, s in enumerate(seeds)]
    tss = [ts[k % 3::3] if k % 2 else ts for k in range(len(plans))]
    tss = [t if p.with_states else None for p, t in zip(plans, tss)]
    many = eng.expand_many(plans, seeds, tss)
    kept = [(r.emitted, r.frontier, r.scanned, None if r.count_only else r.all_rows()) for r in many]
    for k, (p, s, t) in enumerate(zip(plans, seeds, tss)):
        r = eng.expand(p, s, t)
        em, fr, sc, rows = kept[k]
        assert r.emitted == em and r.frontier == fr and r.scanned == sc, k
        if rows is not None:
            one = r.all_rows()
            assert one.keys() == rows.keys()
            for lvl in rows:
                assert np.array_equal(one[lvl], rows[lvl]), (k, lvl)
    eng.close()


def test_double_buffered_async_copy_out(snap_10k):
    """rca_set_double_buffer + rca_result_copy_async: rows keep travelling to the host while the next submissions
    run; a later submission never overwrites rows that are still being copied."""
    import torch
    from k8s_llm_rca_b200 import engine
    node, ev, ns, pod = _kinds(snap_10k, "Node", "Event", "Namespace", "Pod")
    plan = engine.varlen_plan("und", 3, dest_kind=node, excl_kinds=[ev, ns], entity_only=True)
    pods = snap_10k.nodes_of_kind(pod)
    parts = [pods[k::4] for k in range(4)]
    ref_eng = engine.Engine(snap_10k)
    want = [ref_eng.expand(plan, p).rows(3).copy() for p in parts]
    ref_eng.close()
    eng = engine.Engine(snap_10k)
    with pytest.raises(engine.RcaError):
        eng.expand(plan, parts[0]).rows_async(3, np.empty((len(want[0]), 7), np.int32))      # not switched on yet
    eng.set_double_buffer(True)
    bufs = [torch.empty(max(len(w) for w in want) * 7, dtype=torch.int32).pin_memory().numpy() for _ in range(4)]
    views = []
    for rnd in range(3):                                  # arenas alternate: every arena is reused several times
        views.clear()
        for k, p in enumerate(parts):
            r = eng.expand(plan, p)
            views.append(r.rows_async(3, bufs[k]))       # enqueue, do not wait: the next expand() runs meanwhile
        eng.result_wait()
        for k in range(4):
            assert np.array_equal(views[k], want[k]), (rnd, k)
    # the synchronous copy still works with double buffering on, and device pointers of the last TWO results differ
    r1 = eng.expand(plan, parts[0])
    p1 = r1.device_rows(3)[0]
    r2 = eng.expand(plan, parts[1])
    assert np.array_equal(r2.rows(3), want[1]) and r2.device_rows(3)[0] != p1
    eng.set_double_buffer(False)
    assert np.array_equal(eng.expand(plan, parts[2]).rows(3), want[2])
    eng.close()

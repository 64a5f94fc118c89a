"""The oracle is only worth something if it is itself pinned.  The reference holds no golden
vectors for this path (parity unpinned, see oracle/cypher_oracle.py), so the restatement is held
to: (1) its own literal evaluate-the-WHERE-afterwards form, (2) an independent enumeration with
networkx, (3) the C restatement used at scale, (4) the reference's written interval rule."""
import networkx as nx
import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from conftest import rows_to_set
from k8s_llm_rca_b200.props import TableProps, ms_to_ts
from k8s_llm_rca_b200.snapshot import Kind, Snapshot
from oracle.cypher_oracle import PGOracle

KINDS = ["A", "B", "C", "Event", "Namespace"]


def make_graph(n_per_kind, edges):
    """edges: (src_kind, src_local, dst_kind, dst_local, type_index) over KINDS - may hold parallel edges."""
    kinds = [Kind(k, k, "NativeEntity") for k in KINDS]
    start = np.concatenate([[0], np.cumsum(n_per_kind)])
    ets, idx, es, ed, ee = [], {}, [], [], []
    for (sk, sl, dk, dl, t) in edges:
        key = ("R%d" % t, sk, dk, "k%d" % t)
        if key not in idx:
            idx[key] = len(ets)
            ets.append(key)
        es.append(start[sk] + sl % n_per_kind[sk])
        ed.append(start[dk] + dl % n_per_kind[dk])
        ee.append(idx[key])
    snap = Snapshot.build(kinds, n_per_kind, ets, es, ed, ee)
    snap.props = TableProps(snap, [{"kind": KINDS[int(snap.kind_of(v))]} for v in range(snap.n_nodes)])
    return snap


edge_st = st.tuples(st.integers(0, 4), st.integers(0, 3), st.integers(0, 4), st.integers(0, 3), st.integers(0, 1))


@settings(max_examples=60, deadline=None)
@given(st.lists(edge_st, min_size=1, max_size=14), st.booleans(), st.sampled_from([[], ["B"], ["B", "C"]]))
def test_pruned_form_equals_literal_form_and_networkx(edges, directed, inter):
    snap = make_graph([3, 3, 3, 2, 2], edges)
    pg = PGOracle(*snap.to_property_graph())
    lit = pg.varlen_literal("A", "C", inter, directed, max_hops=3)
    pruned = pg.varlen_paths("A", "C", inter, directed, max_hops=3)
    assert lit == pruned
    # independent: simple edge paths of a MultiDiGraph keyed by relationship id, filtered by the same WHERE
    g = nx.MultiDiGraph()
    g.add_nodes_from(range(snap.n_nodes))
    for r in pg.rels:
        g.add_edge(r["start"], r["end"], key=r["id"])
        if not directed and r["start"] != r["end"]:
            g.add_edge(r["end"], r["start"], key=r["id"])
    kind = lambda v: pg.prop(v, "kind")
    want = set()
    for s in [v for v in range(snap.n_nodes) if kind(v) == "A"]:
        for t in [v for v in range(snap.n_nodes) if kind(v) == "C" and v != s]:
            for ep in nx.all_simple_edge_paths(g, s, t, cutoff=3):
                nodes = [ep[0][0]] + [e[1] for e in ep]
                if any(kind(v) in ("Event", "Namespace") for v in nodes):
                    continue
                if inter and not any(kind(v) in inter for v in nodes[1:-1]):
                    continue
                want.add((tuple(nodes), tuple(e[2] for e in ep)))
    assert set(pruned) == want


@settings(max_examples=40, deadline=None)
@given(st.lists(edge_st, min_size=1, max_size=14), st.sampled_from(["out", "in", "und"]), st.integers(1, 4))
def test_c_oracle_equals_python_oracle(edges, direction, hops):
    from oracle.csr import CsrOracle
    snap = make_graph([3, 3, 3, 2, 2], edges)
    pg = PGOracle(*snap.to_property_graph())
    orc = CsrOracle(snap)
    a, c, ev, ns = 0, 2, 3, 4
    if direction == "in":       # the Python oracle walks out/undirected: reverse the graph for 'in'
        rev = PGOracle(pg.nodes, [dict(r, start=r["end"], end=r["start"]) for r in pg.rels])
        want = rev.varlen_paths("A", "C", [], True, max_hops=hops)
    else:
        want = pg.varlen_paths("A", "C", [], direction == "out", max_hops=hops)
    got, _ = orc.varlen(snap.nodes_of_kind(a), direction, hops, dest_kind=c, excl_kinds=[ev, ns])
    assert rows_to_set(got) == want


def test_c_oracle_on_generated_graph(snap_small):
    from oracle.csr import CsrOracle
    snap = snap_small
    pg = PGOracle(*snap.to_property_graph())
    for n in pg.nodes:
        p = n["props"]
        if "kind2" in p or "tag" in p:
            p["ekind"] = p.get("kind2") or p.get("tag")
    ent = lambda r: "ekind" in pg.nodes[r["start"]]["props"] and "ekind" in pg.nodes[r["end"]]["props"]
    orc = CsrOracle(snap)
    K = snap.kind_index
    excl = [K("Event"), K("Namespace")]
    for directed, d in ((True, "out"), (False, "und")):
        for inter_names in ([], ["ReplicaSet", "ConfigMap"]):
            want = pg.varlen_paths("Pod", "Node", inter_names, directed, 3, kind_prop="ekind", rel_ok=ent)
            got, _ = orc.varlen(snap.nodes_of_kind(K("Pod")), d, 3, dest_kind=K("Node"), excl_kinds=excl,
                                inter_kinds=[K(x) for x in inter_names], entity_only=True)
            assert rows_to_set(got) == want
    # src - Namespace - dest, relationships distinct, nodes may coincide (reference :125-129)
    want = pg.via_kind("Pod", "ConfigMap", "Namespace", kind_prop="ekind")
    got, _ = orc.varlen(snap.nodes_of_kind(K("Pod")), "und", 2, min_hops=2, dest_kind=K("ConfigMap"), node_simple=False,
                        pos_kind=[K("Pod"), K("Namespace"), K("ConfigMap")], entity_only=True)
    assert rows_to_set(got) == want and len(want) > 0
    want = pg.single_hop("Pod", "Node", kind_prop="ekind")
    got, _ = orc.varlen(snap.nodes_of_kind(K("Pod")), "und", 1, dest_kind=K("Node"), node_simple=False,
                        pos_kind=[K("Pod"), K("Node")], entity_only=True)
    assert rows_to_set(got) == want


def test_chain_join_semantics(snap_small):
    """Typed MATCH clauses are a natural join on shared aliases; the C chain walk must equal it."""
    from oracle.csr import CsrOracle
    snap = snap_small
    pg = PGOracle(*snap.to_property_graph())
    orc = CsrOracle(snap)
    clauses = [("n1", "Event", "r1", "HasEvent", "evt", "EVENT", "metadata_uid"),
               ("n1", "Event", "r2", "ReferInternal", "n2", "Pod", "involvedObject_uid"),
               ("n2", "Pod", "r3", "ReferInternal", "n3", "PersistentVolumeClaim", "spec_volumes_persistentVolumeClaim_claimName"),
               ("n4", "PersistentVolume", "r4", "ReferInternal", "n3", "PersistentVolumeClaim", "spec_claimRef_uid")]
    seeds = [n["id"] for n in pg.nodes if "EVENT" in n["labels"]]
    rows = pg.chain([{"evt": s} for s in seeds], clauses)
    want = sorted((r["evt"], r["n1"], r["n2"], r["n3"], r["n4"], r["r1"], r["r2"], r["r3"], r["r4"]) for r in rows)
    L = snap.label_index
    ets = [snap.find_etypes("HasEvent", L("Event"), L("EVENT"), "metadata_uid")[0],
           snap.find_etypes("ReferInternal", L("Event"), L("Pod"), "involvedObject_uid")[0],
           snap.find_etypes("ReferInternal", L("Pod"), L("PersistentVolumeClaim"), "spec_volumes_persistentVolumeClaim_claimName")[0],
           snap.find_etypes("ReferInternal", L("PersistentVolume"), L("PersistentVolumeClaim"), "spec_claimRef_uid")[0]]
    got, _ = orc.chain(seeds, ets, [1, 0, 0, 1])
    assert sorted(map(tuple, got.tolist())) == want and len(want) > 0


def test_interval_rule_is_half_open():
    """reference check_state/analyze_root_cause.py:62-68: 5 is in [5, 8) and not in [3, 5)."""
    kinds = [Kind("Pod", "Pod", "NativeEntity"), Kind("Pod", "POD", "NativeState", True)]
    t = lambda x: 1_600_000_000_000 + x * 1000
    snap = Snapshot.build(kinds, [1, 2], [("HasState", 0, 1, "metadata_uid")], [0, 0], [1, 2], [0, 0],
                          [t(3), t(5)], [t(5), t(8)])
    props = [{"id": "p", "kind2": "Pod"}, {"kind": "Pod", "id": "p"}, {"kind": "Pod", "id": "p"}]
    rel = [{"tmin": ms_to_ts(snap.st_tmin[i]), "tmax": ms_to_ts(snap.st_tmax[i])} for i in range(2)]
    snap.props = TableProps(snap, props, rel)
    pg = PGOracle(*snap.to_property_graph())
    from oracle.csr import CsrOracle
    orc = CsrOracle(snap)
    for x, want in ((3, [1]), (4, [1]), (5, [2]), (7, [2]), (8, []), (2, [])):
        assert pg.strict_states("Pod", "p", ms_to_ts(t(x))) == want
        first, cnt, ids = orc.states([0], [t(x)])
        assert [i for i in ids[0].tolist() if i >= 0] == want and cnt[0] == len(want)
    # loose form (:51-60): windows that overlap [tmin, tmax)
    assert pg.loose_states("Pod", "p", ms_to_ts(t(4)), ms_to_ts(t(6))) == [1, 2]
    assert orc.states([0], [t(6)], ts_hi=[t(4)])[2][0][:2].tolist() == [1, 2]


def test_timestamp_strings_order_like_integers():
    from k8s_llm_rca_b200.props import ts_to_ms
    rng = np.random.default_rng(1)
    ms = np.sort(rng.integers(0, 4_000_000_000_000, size=500))
    strs = [ms_to_ts(x) for x in ms]
    assert strs == sorted(strs) and [ts_to_ms(s) for s in strs] == ms.tolist()
    assert ts_to_ms("2020-12-11 06:35:02.011") == 1607668502011          # reference test_check_state.py:41
    with pytest.raises(ValueError):
        ts_to_ms("2020-12-11T06:35:02.011")


def _sorted(a, w=None):
    a = np.asarray(a)
    return a[np.lexsort(a.T[::-1])] if len(a) else a


@pytest.mark.parametrize("pruned", [False, True])
def test_step_functions_equal_the_plain_oracle_plus_state_lookups(snap_small, csr_oracle_lib, pruned):
    """orc_varlen_step / orc_chain_step (the timed CPU baseline: state columns in C, optional schema pruning) give
    the rows of orc_varlen / orc_chain, and their state columns are what orc_states says node by node."""
    from k8s_llm_rca_b200.synth import T_BASE, DAY
    snap = snap_small
    orc = csr_oracle_lib.CsrOracle(snap)
    K = snap.kind_index
    rng = np.random.default_rng(3)
    seeds = snap.nodes_of_kind(K("Pod")).astype(np.int32)
    ts = (T_BASE + rng.integers(-2 * DAY, 33 * DAY, size=len(seeds))).astype(np.int64)
    for direction, hops, dest in (("und", 4, "Node"), ("out", 3, "nfs"), ("und", 3, "ConfigMap"), ("in", 2, "Service")):
        kw = dict(dest_kind=K(dest), excl_kinds=[K("Event"), K("Namespace")], entity_only=True)
        reach = orc.reach(direction, hops, **kw) if pruned else None
        assert (reach is not None) == pruned
        want, sc_plain = orc.varlen(seeds, direction, hops, **kw)
        got, sc, total = orc.varlen_step(seeds, ts, direction, hops, reach=reach, **kw)
        assert total == sum(len(r) for r in want.values()) and got.keys() == want.keys()
        assert sc == sc_plain if not pruned else sc <= sc_plain
        by_seed = dict(zip(seeds.tolist(), ts.tolist()))
        for s, rows in got.items():
            rows = _sorted(rows)
            assert np.array_equal(rows[:, :2 * s + 1], _sorted(want[s]))
            t = np.array([by_seed[v] for v in rows[:, 0].tolist()], np.int64)
            miss = np.zeros(len(rows), np.int64)
            for d in range(s + 1):
                first, cnt, _ = orc.states(rows[:, d], t, limit=1)
                assert np.array_equal(rows[:, 2 * s + 1 + d], first)
                miss |= (cnt == 0).astype(np.int64) << d
            assert np.array_equal(rows[:, 3 * s + 2], miss.astype(np.int32))
        # without timestamps: plain rows
        got2, _, _ = orc.varlen_step(seeds, None, direction, hops, reach=reach, **kw)
        for s in want:
            assert np.array_equal(_sorted(got2[s]), _sorted(want[s]))
    # a plan pruning does not apply to (no destination)
    assert orc.reach("und", 3, dest_kind=-1) is None
    # typed chain with state columns
    L = snap.label_index
    ets = [snap.find_etypes("HasEvent", L("Event"), L("EVENT"), "metadata_uid")[0],
           snap.find_etypes("ReferInternal", L("Event"), L("Pod"), "involvedObject_uid")[0],
           snap.find_etypes("ReferInternal", L("Pod"), L("Node"), "spec_nodeName")[0]]
    ev = snap.nodes_of_kind(L("EVENT")).astype(np.int32)
    ets_ts = (T_BASE + rng.integers(0, 30 * DAY, size=len(ev))).astype(np.int64)
    want, _ = orc.chain(ev, ets, [1, 0, 0])
    got, _ = orc.chain_step(ev, ets_ts, ets, [1, 0, 0])
    assert len(got) == len(want) > 0
    got = _sorted(got)
    assert np.array_equal(got[:, :7], _sorted(want))
    by_seed = dict(zip(ev.tolist(), ets_ts.tolist()))
    t = np.array([by_seed[v] for v in got[:, 0].tolist()], np.int64)
    for d in range(4):
        first, _, _ = orc.states(got[:, d], t, limit=1)
        assert np.array_equal(got[:, 7 + d], first)

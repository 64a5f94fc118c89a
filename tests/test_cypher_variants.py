"""Statements an assistant may write instead of the generated S4 text: one long
pattern, comma-separated patterns, hops without a direction, unnamed nodes and relationships, extra predicates on
any variable, functions, comments, RETURN *.  Each runs text -> QueryRunner (CPU backend of the tests) and is
compared with a brute-force matcher over the plain property graph that shares nothing with the front end."""
import itertools

import pytest

from oracle_backend import oracle_executor

from k8s_llm_rca_b200 import cypher as cy


def brute(nodes, rels, clauses, pred=lambda env: True):
    """clauses: [[(a, label_a, r, types, direction, b, label_b), ...], ...] - one inner list per MATCH clause.
    Returns the list of environments {variable: node id / ('r', rel id)} that satisfy every hop; a relationship is
    bound at most once within one clause (openCypher), freely across clauses."""
    envs = [{}]
    for hops in clauses:
        nxt = []
        for env in envs:
            part = [(dict(env), frozenset())]
            for (a, la, r, types, d, b, lb) in hops:
                grown = []
                for e, used in part:
                    for rel in rels:
                        if rel["type"] not in types or rel["id"] in used:
                            continue
                        ways = {"out": [(rel["start"], rel["end"])], "in": [(rel["end"], rel["start"])],
                                "und": [(rel["start"], rel["end"]), (rel["end"], rel["start"])]}[d]
                        for x, y in dict.fromkeys(ways):
                            if la and la not in nodes[x]["labels"] or lb and lb not in nodes[y]["labels"]:
                                continue
                            if e.get(a, x) != x or e.get(b, y) != y or (r and e.get(r, ("r", rel["id"])) != ("r", rel["id"])):
                                continue
                            e2 = dict(e)
                            e2[a], e2[b] = x, y
                            if r:
                                e2[r] = ("r", rel["id"])
                            grown.append((e2, used | {rel["id"]}))
                part = grown
            nxt.extend(e for e, _ in part)
        envs = nxt
    return [e for e in envs if pred(e)]


def keyed(records):
    out = []
    for rec in records:
        row = []
        for k, v in rec.items():
            row.append((k, ("r", v.id) if type(v).__name__ == "Relationship" else (v.id if hasattr(v, "id") else v)))
        out.append(tuple(sorted(row)))
    return sorted(out)


def expect(envs, columns):
    return sorted(tuple(sorted((c, e[c]) for c in columns)) for e in envs)


@pytest.fixture(scope="module")
def world(snap_small):
    nodes, rels = snap_small.to_property_graph()
    ex = oracle_executor(snap_small)
    # an incident whose Pod mounts a ConfigMap: the S4 text of the reference's template finds one row at least
    for m in snap_small.props.msgs:
        rows = ex.run_query(
            "MATCH (evt:EVENT) WHERE evt.message CONTAINS $m WITH evt LIMIT 1 MATCH (event:Event)-[r1:HasEvent]->(evt) "
            "MATCH (event)-[r2:ReferInternal]->(pod:Pod) WHERE r2.key = 'involvedObject_uid' "
            "MATCH (pod)-[r3:ReferInternal]->(cm:ConfigMap) RETURN pod", {"m": m})
        if rows:
            return snap_small, nodes, rels, ex, m
    raise AssertionError("no incident on a Pod with a ConfigMap in the fixture graph")


def base_clauses(one_clause):
    hops = [("event", "Event", "r1", ["HasEvent"], "out", "evt", "EVENT"),
            ("event", "Event", "r2", ["ReferInternal"], "out", "pod", "Pod"),
            ("pod", "Pod", "r3", ["ReferInternal"], "out", "cm", "ConfigMap")]
    return [hops] if one_clause else [[h] for h in hops]


def test_one_pattern_comma_patterns_and_separate_clauses_agree_with_brute_force(world):
    snap, nodes, rels, ex, m = world
    cols = ["event", "r1", "evt", "r2", "pod", "r3", "cm"]

    def pred(e):
        return (m in nodes[e["evt"]]["props"]["message"] and rels[e["r2"][1]]["props"]["key"] == "involvedObject_uid"
                and rels[e["r3"][1]]["props"]["key"] == "spec_volumes_configMap_name")
    want = expect(brute(nodes, rels, base_clauses(False), pred), cols)
    assert want
    ret = "RETURN event, r1, evt, r2, pod, r3, cm"
    texts = [
        # the template's shape
        f"MATCH (evt:EVENT) WHERE evt.message CONTAINS $m WITH evt LIMIT 1 MATCH (event:Event)-[r1:HasEvent]->(evt) "
        f"MATCH (event)-[r2:ReferInternal]->(pod:Pod) WHERE r2.key = 'involvedObject_uid' "
        f"MATCH (pod)-[r3:ReferInternal]->(cm:ConfigMap) WHERE r3.key = 'spec_volumes_configMap_name' {ret}",
        # one long pattern with a left arrow and inline maps
        f"MATCH (evt:EVENT)<-[r1:HasEvent]-(event:Event)-[r2:ReferInternal {{key: 'involvedObject_uid'}}]->(pod:Pod)"
        f"-[r3:ReferInternal {{key: 'spec_volumes_configMap_name'}}]->(cm:ConfigMap) WHERE evt.message CONTAINS $m {ret}",
        # comma-separated patterns, predicates in any order, a block comment and a trailing semicolon
        f"/* which config map? */ MATCH (event:Event)-[r1:HasEvent]->(evt:EVENT), (event)-[r2:ReferInternal]->(pod:Pod), "
        f"(pod)-[r3:ReferInternal]->(cm:ConfigMap) WHERE r3.key = 'spec_volumes_configMap_name' AND evt.message CONTAINS $m "
        f"AND r2.key = 'involvedObject_uid' {ret};",
        # no directions written at all: the labels decide which way each relationship is stored
        f"MATCH (evt:EVENT)-[r1:HasEvent]-(event:Event)-[r2:ReferInternal]-(pod:Pod)-[r3:ReferInternal]-(cm:ConfigMap) "
        f"WHERE evt.message CONTAINS $m AND r2.key = 'involvedObject_uid' AND r3.key = 'spec_volumes_configMap_name' {ret}",
        # kind tests instead of labels on the far nodes
        f"MATCH (evt:EVENT)<-[r1:HasEvent]-(event:Event)-[r2:ReferInternal]->(pod)-[r3:ReferInternal]->(cm) "
        f"WHERE evt.message CONTAINS $m AND r2.key = 'involvedObject_uid' AND r3.key = 'spec_volumes_configMap_name' "
        f"AND pod.kind2 = 'Pod' AND cm.kind2 = 'ConfigMap' {ret}",
        # RETURN * lists the same variables
        f"MATCH (evt:EVENT)<-[r1:HasEvent]-(event:Event)-[r2:ReferInternal {{key: 'involvedObject_uid'}}]->(pod:Pod)"
        f"-[r3:ReferInternal {{key: 'spec_volumes_configMap_name'}}]->(cm:ConfigMap) WHERE evt.message CONTAINS $m RETURN *",
    ]
    for q in texts:
        assert keyed(ex.run_query(q, {"m": m})) == want, q
    star = ex.run_query(texts[-1], {"m": m})
    assert star[0].keys() == sorted(cols)


def test_extra_predicates_are_applied_to_the_rows(world):
    snap, nodes, rels, ex, m = world
    head = ("MATCH (evt:EVENT) WHERE evt.message CONTAINS $m WITH evt LIMIT 1 MATCH (event:Event)-[r1:HasEvent]->(evt) "
            "MATCH (event)-[r2:ReferInternal]->(pod:Pod) WHERE r2.key = 'involvedObject_uid' MATCH (pod)-[r3:ReferInternal]->(x) ")
    clauses = [[("event", "Event", "r1", ["HasEvent"], "out", "evt", "EVENT")],
               [("event", "Event", "r2", ["ReferInternal"], "out", "pod", "Pod")],
               [("pod", "Pod", "r3", ["ReferInternal"], "out", "x", None)]]
    P, R = (lambda e, v: nodes[e[v]]["props"]), (lambda e, v: rels[e[v][1]]["props"])

    def seeded(e):
        return m in P(e, "evt")["message"] and R(e, "r2")["key"] == "involvedObject_uid"
    cases = [
        ("WHERE r3.key STARTS WITH 'spec_volumes'", lambda e: R(e, "r3")["key"].startswith("spec_volumes")),
        ("WHERE r3.key IN ['spec_nodeName', 'spec_volumes_secret_secretName']",
         lambda e: R(e, "r3")["key"] in ("spec_nodeName", "spec_volumes_secret_secretName")),
        ("WHERE type(r3) = 'ReferInternal' AND x.kind2 <> 'Node'", lambda e: P(e, "x").get("kind2") not in (None, "Node")),
        ("WHERE x.isNative = 'true' AND NOT x.kind2 IN ['Namespace', 'Node']",
         lambda e: P(e, "x").get("isNative") == "true" and P(e, "x").get("kind2") not in (None, "Namespace", "Node")),
        ("WHERE x.name2 IS NOT NULL AND toLower(x.kind2) = 'secret'", lambda e: (P(e, "x").get("kind2") or "").lower() == "secret"),
        ("WHERE x.name2 =~ 'configmap-.*' OR x.kind2 = 'Secret'",
         lambda e: (P(e, "x").get("name2") or "").startswith("configmap-") or P(e, "x").get("kind2") == "Secret"),
        ("WHERE exists(x.no_such_property)", lambda e: False),
        ("WHERE pod.id = x.id", lambda e: False),
        ("WHERE 'Secret' IN labels(x) OR size(r3.key) < 15",
         lambda e: "Secret" in nodes[e["x"]]["labels"] or len(R(e, "r3")["key"]) < 15),
    ]
    seen_rows = 0
    for tail, pred in cases:
        want = expect(brute(nodes, rels, clauses, lambda e: seeded(e) and pred(e)), ["pod", "r3", "x"])
        got = keyed(ex.run_query(head + tail + " RETURN pod, r3, x", {"m": m}))
        assert got == want, tail
        seen_rows += len(want)
    assert seen_rows > 0


def test_unnamed_nodes_and_relationships(world):
    snap, nodes, rels, ex, m = world
    q = ("MATCH (evt:EVENT)<-[:HasEvent]-(:Event)-[:ReferInternal {key: 'involvedObject_uid'}]->(pod:Pod)-[:ReferInternal]->(cm:ConfigMap) "
         "WHERE evt.message CONTAINS $m RETURN *")

    def pred(e):
        return m in nodes[e["evt"]]["props"]["message"] and rels[e["r2"][1]]["props"]["key"] == "involvedObject_uid"
    envs = brute(nodes, rels, base_clauses(True), pred)
    got = ex.run_query(q, {"m": m})
    assert got and got[0].keys() == ["cm", "evt", "pod"]                 # RETURN * names only what the text names
    assert keyed(got) == expect(envs, ["cm", "evt", "pod"])


def test_an_undirected_hop_between_nodes_of_one_kind_matches_both_ways_round():
    """(a:Pod)-[r:Peers]-(b:Pod): each stored relationship answers twice, once per orientation, a self-loop once."""
    import numpy as np
    from k8s_llm_rca_b200.props import TableProps
    from k8s_llm_rca_b200.snapshot import Kind, Snapshot
    kinds = [Kind("Pod", "Pod", "NativeEntity", False)]
    snap = Snapshot.build(kinds, [3], [("Peers", 0, 0, "k")], [0, 1, 2], [1, 2, 2], [0, 0, 0])
    snap.props = TableProps(snap, [{"kind2": "Pod", "name2": f"p{i}"} for i in range(3)])
    ex = oracle_executor(snap)
    got = sorted((r["a"]["name2"], r["b"]["name2"]) for r in ex.run_query("MATCH (a:Pod)-[r:Peers]-(b:Pod) RETURN a, b"))
    assert got == [("p0", "p1"), ("p1", "p0"), ("p1", "p2"), ("p2", "p1"), ("p2", "p2")]
    nodes, rels = snap.to_property_graph()
    want = expect(brute(nodes, rels, [[("a", "Pod", "r", ["Peers"], "und", "b", "Pod")]]), ["a", "b"])
    assert keyed(ex.run_query("MATCH (a:Pod)-[r:Peers]-(b:Pod) RETURN a, b")) == want
    # two undirected hops in ONE clause may not reuse a relationship; in two clauses they may
    one = ex.run_query("MATCH (a:Pod)-[r:Peers]-(b:Pod)-[q:Peers]-(c:Pod) WHERE a.name2 = 'p0' RETURN a, b, c")
    two = ex.run_query("MATCH (a:Pod)-[r:Peers]-(b:Pod) WHERE a.name2 = 'p0' MATCH (b)-[q:Peers]-(c:Pod) RETURN a, b, c")
    assert sorted(r["c"]["name2"] for r in one) == ["p2"]
    assert sorted(r["c"]["name2"] for r in two) == ["p0", "p2"]


def test_predicates_after_the_limit_do_not_choose_the_seed(world):
    """WITH evt LIMIT 1 picks ONE event state; a later test on it filters rows, it does not widen the choice."""
    snap, nodes, rels, ex, m = world
    q = ("MATCH (evt:EVENT) WITH evt LIMIT 1 MATCH (event:Event)-[r1:HasEvent]->(evt) WHERE evt.message CONTAINS $m RETURN evt")
    first_evt = min(v for v in range(len(nodes)) if "EVENT" in nodes[v]["labels"])
    rows = ex.run_query(q, {"m": m})
    assert [r["evt"].id for r in rows] == ([first_evt] if m in nodes[first_evt]["props"]["message"] else [])


@pytest.mark.parametrize("text, fragment", [
    ("MATCH (a:Pod)-[r:ReferInternal]->(b) WHERE c.kind2 = 'x' RETURN a", "`c` not defined"),
    ("MATCH (a:Pod)-[r:ReferInternal]->(b) RETURN count(*)", "aggregation"),
    ("MATCH (a:Pod)-[r:ReferInternal]->(b) RETURN collect(b)", "aggregation"),
    ("MATCH (a:Pod)-[r:ReferInternal]->(b) WHERE nosuchfn(b.x) = 1 RETURN a", "Unknown function"),
    ("OPTIONAL MATCH (a:Pod)-[r:ReferInternal]->(b) RETURN a", "OPTIONAL"),
    ("MATCH (a:Pod)-[r:ReferInternal]->(b) /* never closed RETURN a", "Invalid input"),
])
def test_what_is_not_supported_says_so(snap_small, text, fragment):
    ex = oracle_executor(snap_small)
    with pytest.raises(cy.CypherSyntaxError) as ei:
        ex.run_query(text)
    assert fragment in str(ei.value)


def test_the_reference_statements_against_the_same_brute_force_matcher(snap_small):
    """S2, S3c and S3d as the reference writes them, three ways: text through the runner, the property-graph oracle
    called directly, and the brute-force matcher above - which shares code with neither."""
    from oracle.cypher_oracle import PGOracle
    from test_cypher import S2, S3C, S3D
    nodes, rels = snap_small.to_property_graph()
    ex, pg = oracle_executor(snap_small), PGOracle(nodes, rels)
    types = sorted({r["type"] for r in rels})
    # S2: two clauses that share n1
    for m in snap_small.props.msgs[::5]:
        envs = brute(nodes, rels, [[("n1", "Event", "s1", ["HasEvent"], "out", "N1", "EVENT")],
                                   [("n1", "Event", "r1", ["ReferInternal"], "out", "n2", None)]],
                     lambda e: m in nodes[e["N1"]]["props"]["message"] and rels[e["r1"][1]]["props"]["key"] == "involvedObject_uid")
        want = {nodes[e["n2"]]["props"].get("kind2") for e in envs}
        assert len(want) <= 5                                            # LIMIT 5 cuts nothing here: sets compare
        assert {r["n2.kind2"] for r in ex.run_query(S2, {"message": m})} == want == set(pg.src_kinds(m))
    # S3c / S3d on the schema graph, where the reference runs them (its nodes carry `kind`)
    meta = snap_small.derive_metagraph()
    nodes, rels = meta.to_property_graph()
    ex, pg = oracle_executor(meta), PGOracle(nodes, rels)
    types = sorted({r["type"] for r in rels})
    kind = lambda v: nodes[v]["props"].get("kind")
    rows1 = rows2 = 0
    for src, dst in (("Pod", "Node"), ("Pod", "Pod"), ("PersistentVolume", "PersistentVolumeClaim"), ("Pod", "Service"),
                     ("ConfigMap", "Secret")):
        par = {"srcKind": src, "destKind": dst}
        one = brute(nodes, rels, [[("n1", None, "r1", types, "und", "n2", None)]],
                    lambda e: kind(e["n1"]) == src and kind(e["n2"]) == dst)
        want1 = sorted(((e["n1"], e["n2"]), (e["r1"][1],)) for e in one)
        assert sorted(r["path"].key() for r in ex.run_query(S3C, par)) == want1 == pg.single_hop(src, dst)
        two = brute(nodes, rels, [[("n1", None, "r1", types, "und", "n2", None), ("n2", None, "r2", types, "und", "n3", None)]],
                    lambda e: kind(e["n1"]) == src and kind(e["n2"]) == "Namespace" and kind(e["n3"]) == dst)
        want2 = sorted(((e["n1"], e["n2"], e["n3"]), (e["r1"][1], e["r2"][1])) for e in two)
        assert sorted(r["path"].key() for r in ex.run_query(S3D, par)) == want2 == pg.via_kind(src, dst)
        rows1, rows2 = rows1 + len(want1), rows2 + len(want2)
    assert rows1 and rows2                                                 # each rung had rows in one of the cases at least


def test_more_than_one_predicate_on_the_seed(world):
    """`evt.message CONTAINS .. AND evt.timestamp >= ..`: the text lookup picks the candidates, the rest filters them."""
    snap, nodes, rels, ex, m = world
    evt = next(v for v in range(len(nodes)) if "EVENT" in nodes[v]["labels"] and m in nodes[v]["props"]["message"])
    ts = nodes[evt]["props"]["timestamp"]
    q = ("MATCH (evt:EVENT) WHERE evt.message CONTAINS $m AND evt.timestamp {op} $ts WITH evt LIMIT 1 "
         "MATCH (event:Event)-[r1:HasEvent]->(evt) RETURN evt, event")
    calls = []
    inner = ex.runner.backend.contains
    ex.runner.backend.contains = lambda texts: (calls.append(texts), inner(texts))[1]
    try:
        hit = ex.run_query(q.format(op=">="), {"m": m, "ts": ts})
        miss = ex.run_query(q.format(op=">"), {"m": m, "ts": ts})
    finally:
        ex.runner.backend.contains = inner
    assert [r["evt"].id for r in hit] == [evt] and miss == []
    assert len(calls) == 2                                               # answered through the message arena both times

"""Host logic of the drop-in (front end, runner, types, mirrored stage functions) on the CPU oracle
backend, held to the golden fixture that the REFERENCE's unmodified stage functions produced
(tests/golden/make_golden.py) and - where /root/reference exists - to the reference itself."""
import contextlib
import io
import json
import os
import sys

import numpy as np
import pytest

from conftest import ROOT
from oracle_backend import oracle_executor, record_key

GOLDEN = os.path.join(ROOT, "tests", "golden", "incidents_1k.json")
MIRROR = os.path.join(ROOT, "k8s-llm-rca_b200", "dropin")
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
from make_golden import load_stage_functions, run_incidents  # noqa: E402


@pytest.fixture(scope="module")
def golden():
    return json.load(open(GOLDEN))


@pytest.fixture(scope="module")
def stub_cls():
    from k8s_llm_rca_b200 import dropin
    dropin.install(shim_only=True, force_shim=True)
    import importlib.util
    spec = importlib.util.spec_from_file_location("_stub_assistant", os.path.join(MIRROR, "common", "openai_generic_assistant.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.OpenAIGenericAssistant


def test_mirror_over_oracle_reproduces_golden(golden, snap_1k, stub_cls):
    fm, gq, cs = load_stage_functions(MIRROR)
    assert fm.__file__.startswith(MIRROR)
    got = run_incidents(fm, gq, cs, oracle_executor(snap_1k.derive_metagraph()), oracle_executor(snap_1k), snap_1k,
                        lambda: stub_cls({"summarize the root cause": "{}"}))
    assert got["kinds"] == golden["kinds"]
    assert len(got["incidents"]) == len(golden["incidents"])
    for a, b in zip(got["incidents"], golden["incidents"]):
        assert a["srcKind"] == b["srcKind"] and a["metapaths"] == b["metapaths"]
        assert [x["steps"] for x in a["instances"]] == [x["steps"] for x in b["instances"]]
        assert json.loads(json.dumps([x["rows"] for x in a["instances"]])) == [x["rows"] for x in b["instances"]]
        assert [x["clues"] for x in a["instances"]] == [x["clues"] for x in b["instances"]]


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="the reference tree is only in the build container")
def test_reference_stage_functions_run_unmodified_and_match_golden(golden, snap_1k, stub_cls):
    fm, gq, cs = load_stage_functions("/root/reference")
    assert fm.__file__.startswith("/root/reference") and cs.__file__.startswith("/root/reference")
    got = run_incidents(fm, gq, cs, oracle_executor(snap_1k.derive_metagraph()), oracle_executor(snap_1k), snap_1k,
                        lambda: stub_cls({"summarize the root cause": "{}"}))
    assert json.loads(json.dumps(got["incidents"])) == golden["incidents"]


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="the reference tree is only in the build container")
def test_mirror_emits_the_same_query_semantics_as_the_reference():
    """Same metapath text -> the two generators' queries compile to the same plan."""
    from k8s_llm_rca_b200 import cypher as cy
    mp = ("\n    HasEvent, Event, EVENT, metadata_uid;\n    ReferInternal, Event, Pod, involvedObject_uid;\n"
          "    ReferInternal, Pod, PersistentVolumeClaim, spec_volumes_persistentVolumeClaim_claimName;\n"
          "ReferInternal, PersistentVolume, PersistentVolumeClaim, spec_claimRef_uid;\nUseExternal, PersistentVolume, nfs, spec_nfs_path;\n")
    msg = 'it\'s "quoted"'
    plans = []
    for tree in ("/root/reference", MIRROR):
        _, gq, cs = load_stage_functions(tree)
        with contextlib.redirect_stdout(io.StringIO()):
            q = gq.human_generate_cypher_query(mp, msg)
        p = cy.compile_query(q)
        plans.append((p.seed_var, p.seed_where, p.seed_limit, p.labels, p.walk, [(s.rel_types, s.key) for s in p.steps],
                      [src for (_, _, src) in p.ret.items]))
        s5 = cy.compile_query(cs.find_strict_states("Pod", "x", "2020-12-11 06:35:02.011"))
        l5 = cy.compile_query(cs.find_loose_states("Pod", "x", "2020-12-11 06:35:02.011", "2020-12-11 07:35:02.011"))
        plans[-1] += ((s5.entity_label, s5.state_label, s5.entity_id, s5.lo, s5.hi), (l5.lo, l5.hi))
    assert plans[0] == plans[1]


def test_runner_against_property_graph_oracle(snap_small):
    """Cypher text -> Records over the C oracle must equal the plain-Python property-graph oracle."""
    from oracle.cypher_oracle import PGOracle
    snap = snap_small
    meta = snap.derive_metagraph()
    pg_meta, pg = PGOracle(*meta.to_property_graph()), PGOracle(*snap.to_property_graph())
    mx, sx = oracle_executor(meta), oracle_executor(snap)
    # S1
    recs = mx.run_query("MATCH (n1) WHERE n1.category IN ['NativeEntity', 'ExternalEntity'] RETURN n1.category AS category, n1.kind AS kind")
    assert sorted((r["category"], r["kind"]) for r in recs) == sorted(pg_meta.native_external_kinds())
    # S3a/b/c/d over every pair of kinds
    from test_cypher import S3, S3C, S3D
    kinds = [k.name for k in meta.kinds]
    for src in ("Pod", "PersistentVolumeClaim", "Service", "Event"):
        for dst in kinds:
            for inter in ([], ["PersistentVolume", "ReplicaSet"]):
                par = {"srcKind": src, "destKind": dst, "intermediateKinds": inter}
                got = sorted(r["path"].key() for r in mx.run_query(S3.format(arrow="->"), par))
                assert got == pg_meta.varlen_paths(src, dst, inter, True, 3)
                got = sorted(r["path"].key() for r in mx.run_query(S3.format(arrow="-"), par))
                assert got == pg_meta.varlen_paths(src, dst, inter, False, 3)
            par = {"srcKind": src, "destKind": dst}
            assert sorted(r["path"].key() for r in mx.run_query(S3C, par)) == pg_meta.single_hop(src, dst)
            assert sorted(r["path"].key() for r in mx.run_query(S3D, par)) == pg_meta.via_kind(src, dst)
    # S2 / S5 / S6 on the state graph
    from test_cypher import S2
    for i in (0, 7, len(snap.props.msgs) - 1):
        msg = snap.props.msgs[i]
        assert [r["n2.kind2"] for r in sx.run_query(S2, {"message": msg})] == pg.src_kinds(msg)
    assert sx.run_query(S2, {"message": "no such message"}) == []
    ids, _ = snap.props.messages()
    rng = np.random.default_rng(2)
    from k8s_llm_rca_b200.props import ms_to_ts, node_uuid
    from k8s_llm_rca_b200.synth import T_BASE, DAY
    for v in rng.integers(0, int(snap.kind_start[snap.kind_index("Event")]), size=60).tolist():
        kind = snap.kinds[int(snap.kind_of(v))]
        ts = ms_to_ts(T_BASE + int(rng.integers(-3 * DAY, 33 * DAY)))
        q = (f"MATCH (n1:{kind.label})-[r1:HasState]->(n2:{kind.label.upper()}) WHERE n1.id = '{node_uuid(v)}' "
             f"AND r1.tmin <= '{ts}' AND r1.tmax > '{ts}' RETURN n2 LIMIT 10;")
        assert [r["n2"].id for r in sx.run_query(q)] == pg.strict_states(kind.label, node_uuid(v), ts)
        q6 = f"match (n1:{kind.label}) where n1.id = '{node_uuid(v)}' return n1 limit 1"
        assert [r["n1"].id for r in sx.run_query(q6)] == pg.entity_by_id(kind.label, node_uuid(v))


def test_reference_error_behaviour_is_preserved(snap_small, stub_cls):
    fm, gq, cs = load_stage_functions(MIRROR)
    mx, sx = oracle_executor(snap_small.derive_metagraph()), oracle_executor(snap_small)
    with contextlib.redirect_stdout(io.StringIO()):
        with pytest.raises(IndexError):                       # reference :88, records[0] on no match
            fm.find_srcKind(sx, "a message nobody emitted")
        with pytest.raises(TypeError):                        # reference :133, default argument unusable
            fm.find_metapath(mx, "Pod", "Node")
        with pytest.raises(ValueError):                       # reference :153, min() of nothing
            fm.find_metapath(mx, "ResourceQuota", "image", [])
    from k8s_llm_rca_b200.cypher import CypherSyntaxError
    import neo4j
    assert neo4j.exceptions.CypherSyntaxError is CypherSyntaxError
    with pytest.raises(neo4j.exceptions.CypherSyntaxError):
        sx.run_query("MATCH (n:Pod RETURN n")


def test_types_behave_like_the_driver(snap_small):
    from k8s_llm_rca_b200.graphtypes import Node, Path, Record, Relationship
    n = Node(snap_small, 0)
    assert n["kind2"] == "Pod" and n["no such key"] is None and "id" in n.keys() and n.labels == frozenset({"Pod"})
    r = Relationship(snap_small, 0)
    assert r["kind"] is None and r.type and r["key"] and r.start_node.id == int(snap_small.rel_src[0])
    p = Path([n, r.end_node], [r])
    assert len(p) == 1 and p.nodes[0] == n and p.relationships[0] == r
    rec = Record(["a", "b"], [n, r])
    assert rec["a"] is n and rec[1] is r and len(rec) == 2 and list(rec) == [n, r] and rec.keys() == ["a", "b"]
    with pytest.raises(KeyError):
        rec["missing"]


def test_executor_uri_resolution():
    from k8s_llm_rca_b200 import executor
    with pytest.raises(executor.ServiceUnavailable):
        executor.resolve_snapshot("bolt://10.1.0.174:7687")
    with pytest.raises(executor.ServiceUnavailable):
        executor.resolve_snapshot("snapshot://nothing-registered")
    snap = executor.resolve_snapshot("synth://?objects=200&events=20&seed=5")
    meta = executor.resolve_snapshot("synth://?objects=200&events=20&seed=5#metagraph")
    assert snap.n_nodes > 200 and meta.n_nodes == sum(1 for k in snap.kinds if not k.is_state)
    assert executor.resolve_snapshot("synth://?objects=200&events=20&seed=5") is snap


def test_llm_style_variants_of_the_instance_query(snap_small):
    """Statements an LLM may write instead of the generated template: comma-separated
    patterns in one MATCH, several hops in one pattern, ORDER BY / SKIP / LIMIT - same rows as the template
    form, and Cypher's rule that one MATCH clause binds no relationship twice."""
    snap = snap_small
    sx = oracle_executor(snap)
    msg = snap.props.msgs[3]
    template = f"""MATCH (evt:EVENT) WHERE evt.message CONTAINS {msg!r} WITH evt LIMIT 1
        MATCH (event:Event)-[r1:HasEvent]->(evt) WHERE r1.key = 'metadata_uid'
        MATCH (event)-[r2:ReferInternal]->(pod:Pod) WHERE r2.key = 'involvedObject_uid'
        RETURN event, r1, evt, r2, pod"""
    comma = f"""MATCH (evt:EVENT) WHERE evt.message CONTAINS {msg!r} WITH evt LIMIT 1
        MATCH (event:Event)-[r1:HasEvent {{key:'metadata_uid'}}]->(evt), (event)-[r2:ReferInternal {{key:'involvedObject_uid'}}]->(pod:Pod)
        RETURN event, r1, evt, r2, pod"""
    one_path = f"""MATCH (evt:EVENT) WHERE evt.message CONTAINS {msg!r} WITH evt LIMIT 1
        MATCH (evt)<-[r1:HasEvent]-(event:Event)-[r2:ReferInternal]->(pod:Pod)
        WHERE r1.key = 'metadata_uid' AND r2.key = 'involvedObject_uid'
        RETURN event, r1, evt, r2, pod"""
    want = [tuple(x.id for x in r) for r in sx.run_query(template)]
    assert want
    for q in (comma, one_path):
        assert [tuple(x.id for x in r) for r in sx.run_query(q)] == want
    # one MATCH clause never binds a relationship twice: pods that share a ConfigMap, the pod itself excluded
    pod = want[0][4]
    from k8s_llm_rca_b200.props import node_uuid
    share = ("MATCH (p1:Pod)-[r1:ReferInternal {key:'spec_volumes_configMap_name'}]->(cm:ConfigMap)"
             "<-[r2:ReferInternal {key:'spec_volumes_configMap_name'}]-(p2:Pod) "
             f"WHERE p1.id = '{node_uuid(pod)}' RETURN p1, r1, cm, r2, p2")
    rows = [tuple(x.id for x in r) for r in sx.run_query(share)]
    assert all(r[1] != r[3] for r in rows)
    two = ("MATCH (p1:Pod)-[r1:ReferInternal {key:'spec_volumes_configMap_name'}]->(cm:ConfigMap) "
           f"WHERE p1.id = '{node_uuid(pod)}' "
           "MATCH (cm)<-[r2:ReferInternal {key:'spec_volumes_configMap_name'}]-(p2:Pod) RETURN p1, r1, cm, r2, p2")
    rows2 = [tuple(x.id for x in r) for r in sx.run_query(two)]         # two clauses: the relationship may be reused
    assert sorted(rows) == sorted(r for r in rows2 if r[1] != r[3])
    if rows2:
        assert any(r[1] == r[3] for r in rows2)
    # ORDER BY / SKIP / LIMIT
    q = "MATCH (n:Node) RETURN n.name2 AS name, n ORDER BY name DESC"
    names = [r["name"] for r in sx.run_query(q)]
    assert names == sorted(names, reverse=True) and len(names) == len(snap.nodes_of_kind(snap.kind_index("Node")))
    assert [r["name"] for r in sx.run_query(q + " SKIP 1 LIMIT 2")] == names[1:3]
    ids = [r["n"].id for r in sx.run_query("MATCH (n:Node) RETURN n ORDER BY n.id DESC LIMIT 3")]
    assert ids == sorted(snap.nodes_of_kind(snap.kind_index("Node")).tolist(), reverse=True)[:3]

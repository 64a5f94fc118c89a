"""Every statement shape of the hot path, as TEXT, through the front end and runner (cypher.py, executor.QueryRunner
over the C oracle) against oracle.cypher_oracle.PGOracle called DIRECTLY - the plain-Python statement of the
semantics shares no front end with the runner, so a misreading of a Cypher text by the front end shows up here
(the golden fixture and the GPU battery both pass through the front end and cannot see one).

Shapes: S2, S3a / S3b with and without $intermediateKinds on a graph that has a node
WITHOUT the `kind` property (null in a list predicate is not true), S3c, S3d, S4 with a reversed step and with an
alias bound again by a later clause, S5 / S5' at the boundaries of the half-open interval, S6, DISTINCT ... LIMIT.
The reference's own written examples are literal inputs: the step strings of test_generate_query.py:23-40 and
the worked query of generate_query.py:197-207."""
import contextlib
import io
import os

import numpy as np
import pytest

from conftest import ROOT
from oracle.cypher_oracle import PGOracle
from oracle_backend import oracle_executor
from test_cypher import S2, S3, S3C, S3D, TEMPLATE_EXAMPLE

MIRROR = os.path.join(ROOT, "k8s-llm-rca_b200", "dropin")

# reference test_generate_query.py:23-27 and :34-40 (the second ends without ';', as it does there)
STEPS_SECRET = """
    HasEvent, Event, EVENT, metadata_uid;
    ReferInternal, Event, Pod, involvedObject_uid;
    ReferInternal, Pod, Secret, spec_volumes_secret_secretName;
    """
STEPS_NFS = """
    HasEvent, Event, EVENT, metadata_uid;
    ReferInternal, Event, Pod, involvedObject_uid;
    ReferInternal, Pod, PersistentVolumeClaim, spec_volumes_persistentVolumeClaim_claimName;
    ReferInternal, PersistentVolume, PersistentVolumeClaim, spec_claimRef_uid;
    UseExternal, PersistentVolume, nfs, spec_nfs_path;
    """


def _generators():
    """human_generate_cypher_query of the mirror, and of the reference where its tree exists."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    from make_golden import load_stage_functions
    from k8s_llm_rca_b200 import dropin
    dropin.install(shim_only=True, force_shim=True)
    gens = [("mirror", load_stage_functions(MIRROR)[1].human_generate_cypher_query)]
    if os.path.isdir("/root/reference"):
        gens.append(("reference", load_stage_functions("/root/reference")[1].human_generate_cypher_query))
    return gens


def _clauses(steps_text):
    """(a, L1, r, T, b, L2, key) per step, aliases by kind like the generated query (EVENT -> evt)."""
    steps = [[x.strip() for x in part.strip().split(",")] for part in steps_text.split(";") if part.strip()]
    alias, out = {"EVENT": "evt"}, []
    for i, (t, src, dst, key) in enumerate(steps):
        for k in (src, dst):
            alias.setdefault(k, "n%d" % len(alias))
        out.append((alias[src], src, "r%d" % (i + 1), t, alias[dst], dst, key))
    return out, alias


@pytest.mark.parametrize("steps_text", [STEPS_SECRET, STEPS_NFS])
def test_instance_query_text_against_the_property_graph_oracle(snap_small, steps_text):
    snap = snap_small
    pg = PGOracle(*snap.to_property_graph())
    sx = oracle_executor(snap)
    clauses, alias = _clauses(steps_text)
    checked = 0
    for name, gen in _generators():
        for msg in snap.props.msgs:
            with contextlib.redirect_stdout(io.StringIO()):
                q = gen(steps_text, msg)
            recs = sx.run_query(q)
            seeds = pg.event_seeds(msg)
            want = pg.chain([{"evt": v} for v in seeds], clauses)
            cols = recs[0].keys() if recs else []
            got = sorted(tuple(x.id for x in r) for r in recs)
            exp = sorted(tuple(row[c] for c in cols) for row in want) if recs else []
            assert len(recs) == len(want), (name, msg)
            assert got == exp, (name, msg)
            checked += len(recs)
    assert checked > 0                       # the reversed step (PersistentVolume -> claim, walked from the claim) found rows


def test_worked_example_of_the_template_against_the_oracle(snap_small):
    """generate_query.py:197-207 verbatim, the message replaced by one of this graph's: `(event)` and `(evt)` are
    bound again WITHOUT a label in later clauses."""
    snap = snap_small
    pg = PGOracle(*snap.to_property_graph())
    sx = oracle_executor(snap)
    old = TEMPLATE_EXAMPLE.split("CONTAINS ")[1].split("\n")[0]
    hits = 0
    for msg in snap.props.msgs:
        q = TEMPLATE_EXAMPLE.replace(old, repr(msg))
        recs = sx.run_query(q)
        want = pg.chain([{"evt": v} for v in pg.event_seeds(msg)],
                        [("event", "Event", "r1", "HasEvent", "evt", None, "metadata_uid"),
                         ("event", None, "r2", "ReferInternal", "pod", "Pod", "involvedObject_uid"),
                         ("pod", None, "r3", "ReferInternal", "configMap", "ConfigMap", "spec_volumes_configMap_name")])
        cols = ["event", "r1", "evt", "r2", "pod", "r3", "configMap"]
        assert sorted(tuple(x.id for x in r) for r in recs) == sorted(tuple(row[c] for c in cols) for row in want)
        assert not recs or recs[0].keys() == cols
        hits += len(recs)
    assert hits > 0


def _graph_with_a_kindless_node():
    """A-B-C chains where one middle node has NO `kind` property (label Hidden)."""
    from k8s_llm_rca_b200 import ingest
    lines = []
    kinds = {"a": "A", "b": "B", "c": "C", "h": "Hidden", "e": "Event"}
    ids = {}
    for name in ("a0", "a1", "b0", "b1", "h0", "c0", "c1", "e0"):
        ids[name] = str(len(ids))
        props = {} if name[0] == "h" else {"kind": kinds[name[0]]}
        lines.append({"type": "node", "id": ids[name], "labels": [kinds[name[0]]], "properties": props})
    edges = [("a0", "b0"), ("b0", "c0"), ("a0", "h0"), ("h0", "c0"), ("a1", "b1"), ("b1", "c1"), ("c0", "a1"), ("a1", "e0"),
             ("e0", "c1"), ("a0", "c1"), ("a0", "c1"), ("h0", "b1")]
    for i, (u, v) in enumerate(edges):
        lines.append({"type": "relationship", "id": str(i), "label": "R", "start": {"id": ids[u]}, "end": {"id": ids[v]},
                      "properties": {"key": "k%d" % (i % 2)}})
    return ingest.load_dump(lines)


def test_varlen_text_on_a_graph_with_a_node_without_kind():
    snap = _graph_with_a_kindless_node()
    pg = PGOracle(*snap.to_property_graph())
    ex = oracle_executor(snap)
    assert any(pg.prop(v, "kind") is None for v in range(snap.n_nodes))
    n = 0
    for src in ("A", "B", "C"):
        for dst in ("A", "B", "C"):
            for inter in ([], ["B"], ["Hidden"], ["B", "C"], None):
                par = {"srcKind": src, "destKind": dst, "intermediateKinds": inter}
                for arrow, directed in (("->", True), ("-", False)):
                    got = sorted(r["path"].key() for r in ex.run_query(S3.format(arrow=arrow), par))
                    assert got == pg.varlen_literal(src, dst, inter, directed, 3), (src, dst, inter, arrow)
                    assert got == pg.varlen_paths(src, dst, inter, directed, 3)
                    n += len(got)
            par = {"srcKind": src, "destKind": dst}
            assert sorted(r["path"].key() for r in ex.run_query(S3C, par)) == pg.single_hop(src, dst)
            assert sorted(r["path"].key() for r in ex.run_query(S3D, par)) == pg.via_kind(src, dst)
    assert n > 20
    # parallel relationships are distinct paths; a path through the kindless node is never returned
    got = [r["path"] for r in ex.run_query(S3.format(arrow="->"), {"srcKind": "A", "destKind": "C", "intermediateKinds": []})]
    assert sum(1 for p in got if len(p) == 1) == 2
    hidden = [v for v in range(snap.n_nodes) if pg.prop(v, "kind") is None][0]
    assert all(hidden not in [x.id for x in p.nodes] for p in got)


def test_state_queries_at_interval_boundaries(snap_small):
    """S5 strict and S5' loose, text -> records, at tmin (inside), tmax (outside), one millisecond either side."""
    from k8s_llm_rca_b200.props import ms_to_ts, node_uuid
    snap = snap_small
    pg = PGOracle(*snap.to_property_graph())
    sx = oracle_executor(snap)
    n = 0
    for v in range(0, int(snap.kind_start[snap.kind_index("Event")]), 7):
        lo, hi = int(snap.st_off[v]), int(snap.st_off[v + 1])
        if lo == hi:
            continue
        kind = snap.kinds[int(snap.kind_of(v))]
        uid = node_uuid(v)
        for j in range(lo, hi):
            for t in (snap.st_tmin[j] - 1, snap.st_tmin[j], snap.st_tmax[j] - 1, snap.st_tmax[j]):
                if not 0 <= int(t) < 250_000_000_000_000:          # open-ended intervals end at the year 9999: no 'hour later'
                    continue
                ts = ms_to_ts(int(t))
                q = (f"MATCH (n1:{kind.label})-[r1:HasState]->(n2:{kind.label.upper()}) WHERE n1.id = '{uid}' "
                     f"AND r1.tmin <= '{ts}' AND r1.tmax > '{ts}' RETURN n2 LIMIT 10;")
                assert [r["n2"].id for r in sx.run_query(q)] == pg.strict_states(kind.label, uid, ts)
                t2 = ms_to_ts(int(t) + 3_600_000)
                q2 = (f"MATCH (n1:{kind.label})-[r1:HasState]->(n2:{kind.label.upper()}) WHERE n1.id = '{uid}' "
                      f"AND r1.tmin <= '{t2}' AND r1.tmax > '{ts}' RETURN n2 LIMIT 10;")
                assert [r["n2"].id for r in sx.run_query(q2)] == pg.loose_states(kind.label, uid, ts, t2)
                n += 1
    assert n > 50


def test_source_kind_statement_distinct_limit(snap_small):
    """S2: DISTINCT n2.kind2 LIMIT 5 - for every message of the graph, and for a fragment several messages share."""
    snap = snap_small
    pg = PGOracle(*snap.to_property_graph())
    sx = oracle_executor(snap)
    for msg in list(snap.props.msgs) + ["not found", "e", "no such text anywhere"]:
        got = [r["n2.kind2"] for r in sx.run_query(S2, {"message": msg})]
        want = pg.src_kinds(msg)
        assert sorted(got) == sorted(want) if len(want) < 5 else (len(got) == 5 and set(got) <= set(pg.src_kinds(msg, limit=99)))
    q6 = "match (n1:{k})\nwhere n1.id = '{i}'\nreturn n1\nlimit 1"
    from k8s_llm_rca_b200.props import node_uuid
    for v in np.random.default_rng(1).integers(0, int(snap.kind_start[snap.kind_index("Event")]), size=40).tolist():
        k = snap.kinds[int(snap.kind_of(v))].label
        assert [r["n1"].id for r in sx.run_query(q6.format(k=k, i=node_uuid(v)))] == pg.entity_by_id(k, node_uuid(v))
        assert sx.run_query(q6.format(k=k, i="no-such-id")) == []

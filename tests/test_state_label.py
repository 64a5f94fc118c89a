"""The state query applies the label of the state node BEFORE its LIMIT 10 (reference
check_state/analyze_root_cause.py:70-79: `(n1:Kind)-[r1:HasState]->(n2:KIND) ... RETURN n2 LIMIT 10`).

A hand-made dump: one Pod with 16 HasState intervals that all hold the timestamp, alternating between POD and
NODE state nodes (a shape the generator never produces), and a second Pod that shares the first one's `id`.
Truncating to ten ids first and filtering by label afterwards would return five rows; the statement returns eight
per entity.  Checked on the oracle backend (CPU) and through the C ABI (GPU) against oracle.cypher_oracle.PGOracle
called directly."""
import pytest


def _dump():
    lines = []

    def node(i, label, **props):
        lines.append({"type": "node", "id": str(i), "labels": [label], "properties": props})

    node(0, "Pod", kind2="Pod", isNative="true", id="pod-a", name2="pod-0")
    node(1, "Pod", kind2="Pod", isNative="true", id="pod-a", name2="pod-1")      # same id on purpose
    node(2, "Node", kind2="Node", isNative="true", id="node-a", name2="node-0")
    nid, rid = 3, 0
    for owner in (0, 1):
        for j in range(16):
            label = "POD" if j % 2 == 0 else "NODE"
            node(nid, label, kind=label.capitalize(), id=f"st-{nid}")
            lines.append({"type": "relationship", "id": str(rid), "label": "HasState", "start": {"id": str(owner)},
                          "end": {"id": str(nid)}, "properties": {"tmin": "2024-01-01 00:00:00.000",
                                                                  "tmax": "2024-02-01 00:00:00.000"}})
            nid += 1
            rid += 1
    lines.append({"type": "relationship", "id": str(rid), "label": "ReferInternal", "start": {"id": "0"}, "end": {"id": "2"},
                  "properties": {"key": "spec_nodeName"}})
    return lines


Q = ("MATCH (n1:Pod)-[r1:HasState]->(n2:POD) WHERE n1.id = 'pod-a' AND r1.tmin <= '{ts}' AND r1.tmax > '{ts}' "
     "RETURN n2 LIMIT 10;")


def _check(executor, snap):
    from oracle.cypher_oracle import PGOracle
    pg = PGOracle(*snap.to_property_graph())
    ts = "2024-01-15 12:00:00.000"
    want = pg.strict_states("Pod", "pod-a", ts)
    assert len(want) == 10                            # 16 POD states over the two entities, cut at ten
    got = [r["n2"].id for r in executor.run_query(Q.format(ts=ts))]
    assert got == want
    assert all(r["n2"]["kind"] == "Pod" for r in executor.run_query(Q.format(ts=ts)))
    # the boundary of the half-open interval (analyze_root_cause.py:62-68)
    assert executor.run_query(Q.format(ts="2024-02-01 00:00:00.000")) == []
    assert len(executor.run_query(Q.format(ts="2024-01-01 00:00:00.000"))) == 10
    # entity variable: every row carries the entity its state belongs to
    q2 = Q.replace("RETURN n2 LIMIT 10", "RETURN n1, n2 LIMIT 40").format(ts=ts)
    rows = executor.run_query(q2)
    assert len(rows) == 16
    owners = {}
    for r in pg.rels:
        if r["type"] == "HasState":
            owners[r["end"]] = r["start"]
    assert all(owners[r["n2"].id] == r["n1"].id for r in rows)


def test_label_before_limit_on_the_oracle_backend():
    from k8s_llm_rca_b200 import ingest
    from oracle_backend import oracle_executor
    snap = ingest.load_dump(_dump())
    _check(oracle_executor(snap), snap)


@pytest.mark.gpu
def test_label_before_limit_through_the_c_abi():
    from k8s_llm_rca_b200 import executor, ingest
    snap = ingest.load_dump(_dump())
    executor.register_snapshot("state-label", snap)
    ex = executor.GpuQueryExecutor("snapshot://state-label")
    try:
        _check(ex, snap)
    finally:
        ex.close()

"""The drop-in executor on the GPU: Cypher text in, Records out, through the C ABI - held to the
golden fixture (produced by the reference's unmodified stage functions over the oracle) and,
query by query, to the oracle-backed runner."""
import json
import os
import sys

import numpy as np
import pytest

from conftest import ROOT
from oracle_backend import oracle_executor, record_key

pytestmark = pytest.mark.gpu
MIRROR = os.path.join(ROOT, "k8s-llm-rca_b200", "dropin")
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
from make_golden import load_stage_functions, run_incidents  # noqa: E402


@pytest.fixture(scope="module")
def executors(snap_1k):
    from k8s_llm_rca_b200 import dropin, executor
    dropin.install(shim_only=True, force_shim=True)
    executor.register_snapshot("config1", snap_1k)
    state = executor.GpuQueryExecutor("snapshot://config1", "neo4j", "ignored")
    meta = executor.GpuQueryExecutor("snapshot://config1#metagraph", "neo4j", "ignored")
    yield meta, state
    meta.close()
    state.close()


def _stub():
    import importlib.util
    spec = importlib.util.spec_from_file_location("_stub_assistant", os.path.join(MIRROR, "common", "openai_generic_assistant.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.OpenAIGenericAssistant


def test_incident_flow_on_gpu_reproduces_golden(executors, snap_1k):
    golden = json.load(open(os.path.join(ROOT, "tests", "golden", "incidents_1k.json")))
    meta, state = executors
    fm, gq, cs = load_stage_functions(MIRROR)
    Stub = _stub()
    got = run_incidents(fm, gq, cs, meta, state, snap_1k, lambda: Stub({"summarize the root cause": "{}"}))
    assert got["kinds"] == golden["kinds"]
    assert json.loads(json.dumps(got["incidents"])) == golden["incidents"]


def test_neo4j_shim_driver_runs_on_the_gpu(snap_1k):
    """The path the reference's own common/neo4j_query_executor.py takes: GraphDatabase.driver(...)."""
    import neo4j
    from k8s_llm_rca_b200 import executor
    executor.register_snapshot("config1", snap_1k)
    drv = neo4j.GraphDatabase.driver("snapshot://config1#metagraph", auth=("neo4j", "x"))
    drv.verify_connectivity()
    with drv.session() as session:
        recs = list(session.run("MATCH (n1) WHERE n1.category IN ['NativeEntity'] RETURN n1.kind AS kind"))
    assert "Pod" in [r["kind"] for r in recs]
    drv.close()
    bad = neo4j.GraphDatabase.driver("bolt://10.1.0.174:7687", auth=("neo4j", "x"))
    with pytest.raises(neo4j.exceptions.ServiceUnavailable):
        bad.verify_connectivity()


def test_query_battery_gpu_equals_oracle(executors, snap_1k):
    from test_cypher import S2, S3, S3C, S3D
    from k8s_llm_rca_b200.props import ms_to_ts, node_uuid
    from k8s_llm_rca_b200.synth import DAY, T_BASE
    meta, state = executors
    ometa, ostate = oracle_executor(snap_1k.derive_metagraph()), oracle_executor(snap_1k)
    same = lambda a, b: sorted(map(record_key, a)) == sorted(map(record_key, b))
    kinds = [k.name for k in snap_1k.kinds if not k.is_state]
    n = 0
    for src in kinds:
        for dst in kinds:
            for inter in ([], ["PersistentVolume", "Pod"]):
                par = {"srcKind": src, "destKind": dst, "intermediateKinds": inter}
                for arrow in ("->", "-"):
                    q = S3.format(arrow=arrow)
                    a, b = meta.run_query(q, par), ometa.run_query(q, par)
                    assert same(a, b)
                    n += len(a)
            par = {"srcKind": src, "destKind": dst}
            assert same(meta.run_query(S3C, par), ometa.run_query(S3C, par))
            assert same(meta.run_query(S3D, par), ometa.run_query(S3D, par))
    assert n > 100
    # longer hop bounds than the reference's literal 3 (BASELINE configs 2 and 4)
    for hops in (4, 5):
        q = S3.format(arrow="-").replace("*1..3", "*1..%d" % hops)
        par = {"srcKind": "Pod", "destKind": "nfs", "intermediateKinds": []}
        assert same(meta.run_query(q, par), ometa.run_query(q, par))
    # S2 and the state queries on the instance graph
    for i in range(0, len(snap_1k.props.msgs), 9):
        par = {"message": snap_1k.props.msgs[i]}
        assert same(state.run_query(S2, par), ostate.run_query(S2, par))
    rng = np.random.default_rng(4)
    hit = miss = 0
    for v in rng.integers(0, int(snap_1k.kind_start[snap_1k.kind_index("Event")]), size=150).tolist():
        kind = snap_1k.kinds[int(snap_1k.kind_of(v))]
        ts = ms_to_ts(T_BASE + int(rng.integers(-3 * DAY, 33 * DAY)))
        ts2 = ms_to_ts(T_BASE + int(rng.integers(33 * DAY, 40 * DAY)))
        for lo, hi in ((ts, ts), (ts2, ts)):
            q = (f"MATCH (n1:{kind.label})-[r1:HasState]->(n2:{kind.label.upper()}) WHERE n1.id = '{node_uuid(v)}' "
                 f"AND r1.tmin <= '{lo}' AND r1.tmax > '{hi}' RETURN n2 LIMIT 10;")
            a, b = state.run_query(q), ostate.run_query(q)
            assert same(a, b)
            hit += bool(a)
            miss += not a
    assert hit and miss
    with pytest.raises(Exception) as ei:
        state.run_query("MATCH (n:Pod RETURN n")
    assert type(ei.value).__name__ == "CypherSyntaxError"


def test_assistant_written_variants_of_the_instance_query_on_the_gpu(executors, snap_1k):
    """The shapes of tests/test_cypher_variants.py (one pattern, no directions, unnamed parts, extra predicates,
    RETURN *) through the GPU executor, row for row against the oracle-backed runner."""
    _, state = executors
    ostate = oracle_executor(snap_1k)

    def same(a, b):
        return sorted(map(repr, map(record_key, a))) == sorted(map(repr, map(record_key, b))) and \
            [r.keys() for r in a[:1]] == [r.keys() for r in b[:1]]
    tails = [
        "MATCH (evt:EVENT)<-[r1:HasEvent]-(event:Event)-[r2:ReferInternal {key: 'involvedObject_uid'}]->(pod:Pod)"
        "-[r3:ReferInternal]->(x) WHERE evt.message CONTAINS $m RETURN *",
        "MATCH (evt:EVENT)-[r1:HasEvent]-(event:Event)-[r2:ReferInternal]-(pod:Pod)-[r3:ReferInternal]-(x:ConfigMap) "
        "WHERE evt.message CONTAINS $m AND r2.key = 'involvedObject_uid' RETURN event, r1, evt, r2, pod, r3, x",
        "MATCH (evt:EVENT)<-[:HasEvent]-(:Event)-[:ReferInternal {key: 'involvedObject_uid'}]->(pod:Pod)-[r3:ReferInternal]->(x) "
        "WHERE evt.message CONTAINS $m AND x.isNative = 'true' AND NOT x.kind2 IN ['Namespace', 'Node'] RETURN pod, r3, x",
        "/* claims and what binds them */ MATCH (evt:EVENT) WHERE evt.message CONTAINS $m WITH evt LIMIT 1 "
        "MATCH (event:Event)-[r1:HasEvent]->(evt), (event)-[r2:ReferInternal]->(pod:Pod), (pod)-[r3:ReferInternal]->(c), (pv)-[r4:ReferInternal]->(c) "
        "WHERE r2.key = 'involvedObject_uid' AND c.kind2 = 'PersistentVolumeClaim' AND type(r4) = 'ReferInternal' AND pv.kind2 = 'PersistentVolume' "
        "RETURN pod, c, pv, r4",
    ]
    rows = 0
    for i in range(0, len(snap_1k.props.msgs), 7):
        par = {"m": snap_1k.props.msgs[i]}
        for q in tails:
            a, b = state.run_query(q, par), ostate.run_query(q, par)
            assert same(a, b), q
            rows += len(a)
    assert rows > 50

"""Incident replay on the GPU: the batched path equals the one-by-one path (the reference's call
order through run_query), and both equal the golden fixture."""
import json
import os

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def test_batched_replay_equals_one_by_one_and_golden(snap_1k):
    from k8s_llm_rca_b200 import executor, replay
    executor.register_snapshot("config1", snap_1k)
    sx = executor.GpuQueryExecutor("snapshot://config1")
    mx = executor.GpuQueryExecutor("snapshot://config1#metagraph")
    rp = replay.IncidentReplayer(sx, mx)
    msgs = list(snap_1k.props.msgs)
    batch = rp.run_batch(msgs + ["a message no event carries"] + msgs[:5])
    assert batch[len(msgs)] == {"srcKind": None, "metapaths": [], "rows": [], "missing": []}
    some_missing = 0
    for i, m in enumerate(msgs):
        one = rp.run_one(m)
        assert batch[i]["srcKind"] == one["srcKind"]
        assert batch[i]["metapaths"] == one["metapaths"]
        assert batch[i]["rows"] == one["rows"], (i, m)
        assert batch[i]["missing"] == one["missing"]
        some_missing += any(x for x in one["missing"])
    assert some_missing > 0                                   # the 'state missing' branch is exercised
    assert batch[len(msgs) + 1:] == batch[:5]                 # repeated incidents give repeated answers
    golden = json.load(open(os.path.join(ROOT, "tests", "golden", "incidents_1k.json")))
    for inc in golden["incidents"]:
        got = batch[inc["event"]]
        assert got["srcKind"] == inc["srcKind"]
        want_rows = sorted(tuple(x[1] for x in row) for ins in inc["instances"] for row in ins["rows"])
        assert got["rows"] == want_rows
    prof = replay.latency_profile(rp, msgs, n_single=20)
    assert prof["single_p50_ms"] > 0 and prof["batch_ms_per_incident"] < prof["single_p50_ms"]
    sx.close()
    mx.close()

"""The end-to-end incident driver, offline: the reference's own test_all.py main()
where its tree exists, and the mirror driver (dropin/test_all.py) everywhere, run to completion against a
snapshot with recorded assistant answers - k8s_llm_rca_b200.l3.run_driver.  The recorded answers take the run
through both retry loops of the driver (reference test_all.py:63-83 JSON errors, :99-122 Cypher errors) and
through its fall-back to the deterministic template (:127-131).

tests/golden/l3_trace.json is the transcript shape of the REFERENCE's driver (tests/golden/make_l3_trace.py);
the mirror must produce the same shape, over the CPU oracle backend and over the CUDA path."""
import json
import os

import pytest

from conftest import ROOT

FIXTURE = os.path.join(ROOT, "tests", "golden", "l3_trace.json")
MIRROR = os.path.join(ROOT, "k8s-llm-rca_b200", "dropin")


@pytest.fixture(scope="module")
def trace():
    return json.load(open(FIXTURE))


def _assert_ladders(shape):
    text = "\n".join(shape)
    assert shape.count("test find destKind and relevantResources") == 2 and "JSON Error occurred" in shape   # :63-83
    assert [s for s in shape if s.startswith("attempt")] == ["attempt = 0", "attempt = 1", "attempt = 2"]        # :99-122
    assert "An unexpected error occurred" in shape and "Cypher Syntax Error occurred" in shape
    assert text.index("attempt = 2") < text.index("manually generate cypher query") < text.index("entity Pod")  # :127-131
    assert [s.split()[1] for s in shape if s.startswith("entity")] == ["Pod", "PersistentVolumeClaim", "PersistentVolume", "nfs"]
    assert shape[0] == "create executor and init connection" and shape[-1] == "close connection"


def test_fixture_has_both_retry_ladders_and_the_fallback(trace):
    _assert_ladders(trace["shape"])


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="the reference tree is only in the build container")
def test_reference_driver_runs_unmodified_to_completion(trace):
    from k8s_llm_rca_b200 import l3
    from oracle_backend import oracle_executor_factory
    snap, _ = l3.incident_snapshot(trace["message"])
    text, _ = l3.run_driver("/root/reference", snap, trace["recorded"], executor_factory=oracle_executor_factory)
    assert l3.transcript_shape(text) == trace["shape"]
    assert "The code started at" in text                              # the reference's own closing lines were reached


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="the reference tree is only in the build container")
def test_reference_file_driver_runs_unmodified_and_writes_its_result(trace, tmp_path):
    """The reference's second driver, test_with_file.py:20-225: messages from ./data/*.csv (it runs the second row), one
    JSON result per message appended to ./output/*.json with attempt counts, queries, reports and token usage."""
    from k8s_llm_rca_b200 import l3
    from oracle_backend import oracle_executor_factory
    snap, _ = l3.incident_snapshot(trace["message"])
    out_file = l3.with_file_workdir(str(tmp_path), ["a message no event of this graph carries", trace["message"], "a third row"])
    text, _ = l3.run_driver("/root/reference", snap, trace["recorded"], script="test_with_file.py",
                            executor_factory=oracle_executor_factory, workdir=str(tmp_path))
    shape = [s for s in l3.transcript_shape(text)]
    assert shape == trace["shape"]                            # the same ladders as test_all.py on the same incident
    result = json.loads(open(out_file).read())
    assert result["error_message"] == trace["message"] and result["locator_attempts"] == 2
    (analysis,) = result["analysis"]
    assert analysis["cypher_attempts"] == 3 and "human_cypher_query" in analysis and "MATCH (evt:EVENT)" in analysis["human_cypher_query"]
    assert len(analysis["statepath"]) == 1 and json.loads(analysis["statepath"][0]["report"])["conclusion"] == "recorded"
    assert set(analysis["statepath"][0]["clue"]) == {s.split(" ", 1)[1].replace(" ", "(") + ")" for s in trace["shape"] if s.startswith("entity")}
    assert set(result["token_usage"]) == {"prompt_tokens", "completion_tokens", "total_tokens"} and result["time_cost"] >= 0


def test_mirror_driver_over_oracle_backend_matches_the_reference_transcript(trace):
    from k8s_llm_rca_b200 import l3
    from oracle_backend import oracle_executor_factory
    snap, evt = l3.incident_snapshot(trace["message"])
    text, results = l3.run_driver(MIRROR, snap, trace["recorded"], main_kwargs={"messages": [trace["message"]]},
                                  executor_factory=oracle_executor_factory)
    assert l3.transcript_shape(text) == trace["shape"]
    assert len(results) == 1 and results[0][1][0] == evt              # one instance row, starting at the incident's EVENT
    assert json.loads(results[0][2])["conclusion"] == "recorded"
    assert set(results[0][3]) == {s.split(" ", 1)[1].replace(" ", "(") + ")" for s in trace["shape"] if s.startswith("entity")}


def test_uri_alias_and_unknown_uri(snap_small):
    from k8s_llm_rca_b200 import executor
    executor.register_snapshot("alias-test", snap_small)
    executor.register_uri_alias("bolt://10.9.9.9:7687", "snapshot://alias-test#metagraph")
    assert executor.resolve_snapshot("bolt://10.9.9.9:7687").n_nodes == snap_small.derive_metagraph().n_nodes
    with pytest.raises(executor.ServiceUnavailable):
        executor.resolve_snapshot("bolt://10.9.9.10:7687")


@pytest.mark.gpu
def test_mirror_driver_on_the_gpu_matches_the_reference_transcript(trace):
    from k8s_llm_rca_b200 import l3
    snap, evt = l3.incident_snapshot(trace["message"])
    text, results = l3.run_driver(MIRROR, snap, trace["recorded"], main_kwargs={"messages": [trace["message"]]})
    assert l3.transcript_shape(text) == trace["shape"]
    assert len(results) == 1 and results[0][1][0] == evt

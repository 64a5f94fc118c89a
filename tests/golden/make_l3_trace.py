#!/usr/bin/env python
"""Generates tests/golden/l3_trace.json: the transcript shape of the REFERENCE's own end-to-end driver
(/root/reference/test_all.py, main() called unmodified) run offline over the CPU oracle backend with recorded
assistant answers - k8s_llm_rca_b200.l3.run_driver.  The recorded answers make the run take both of the
driver's retry loops (a malformed JSON answer; an answer without a code block, then a syntactically wrong
query, then a query that returns nothing) and its fall-back to the deterministic template.

Run in this container only (needs /root/reference):   python tests/golden/make_l3_trace.py
The `-m "not gpu"` tests hold the mirror driver over the oracle backend to this fixture, the `-m gpu` tests the
mirror driver over the CUDA path; /root/reference is never read at test time.
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

# the incident the reference's driver runs: errorMessages[1:2] of its hard-wired list (test_all.py:40-54)
MESSAGE_FROM = ("/root/reference/test_all.py", 1)

MALFORMED_JSON = "Here is my analysis.\n```json\n{'DestinationKind': nfs, RelevantResources: [Pod, nfs]}\n```"
GOOD_JSON = ('```json\n{"SourceKind": "Pod", "DestinationKind": "nfs", "RelevantResources": ["Pod", '
             '"PersistentVolumeClaim", "PersistentVolume", "nfs"], "PrimaryPath": []}\n```')
NO_CODE_BLOCK = "MATCH (evt:EVENT) RETURN evt   (the answer forgot the code fence)"
BAD_SYNTAX = "```cypher\nMATCH (evt:EVENT WHERE evt.message CONTAINS 'MountVolume' RETURN evt\n```"
NO_ROWS = ("```cypher\nMATCH (evt:EVENT) WHERE evt.message CONTAINS 'MountVolume.SetUp failed' WITH evt LIMIT 1\n"
           "MATCH (e:Event)-[r1:HasEvent]->(evt:EVENT) WHERE r1.key = 'metadata_uid'\n"
           "MATCH (e:Event)-[r2:ReferInternal]->(p:Pod) WHERE r2.key = 'involvedObject_name'\n"
           "RETURN evt, r1, e, r2, p\n```")
RECORDED = {
    "k8s-root-cause-locator": {"": [MALFORMED_JSON, GOOD_JSON]},
    "cypher-query-generator": {"the provided metapath is": [NO_CODE_BLOCK, BAD_SYNTAX, NO_ROWS]},
    "k8s-state-semantic-analyzer": {"summarize the root cause": '{"summary": [], "conclusion": "recorded", "resolution": "recorded"}',
                                    "JSON": "recorded: nothing in this state explains the error"},
}


def reference_message():
    """The message string out of the reference's source, by evaluating its list literal (not by importing the script)."""
    import ast
    path, index = MESSAGE_FROM
    tree = ast.parse(open(path).read())
    for node in ast.walk(tree):
        if isinstance(node, ast.Assign) and getattr(node.targets[0], "id", "") == "errorMessages":
            return ast.literal_eval(node.value)[index]
    raise SystemExit("errorMessages not found in " + path)


def main():
    if not os.path.isdir("/root/reference"):
        raise SystemExit("needs /root/reference (run in the build container)")
    from k8s_llm_rca_b200 import l3
    from oracle_backend import oracle_executor_factory
    message = reference_message()
    snap, _ = l3.incident_snapshot(message)
    text, _ = l3.run_driver("/root/reference", snap, RECORDED, executor_factory=oracle_executor_factory)
    out = {"message": message, "recorded": RECORDED, "shape": l3.transcript_shape(text),
           "source": "/root/reference/test_all.py main(), unmodified, over oracle/csr_oracle.c; snapshot = l3.incident_snapshot(message)"}
    with open(os.path.join(HERE, "l3_trace.json"), "w") as f:
        json.dump(out, f, indent=1)
    print("\n".join(out["shape"]))


if __name__ == "__main__":
    main()

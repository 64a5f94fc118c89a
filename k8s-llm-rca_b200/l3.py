"""Running an end-to-end incident driver offline.

The reference's drivers (test_all.py:18-157, test_with_file.py) open two Neo4j servers at hard-wired bolt
addresses and three OpenAI assistants.  `run_driver` lets such a script run to completion with neither:
  * the bolt addresses become aliases of a snapshot (executor.register_uri_alias) and the module importable
    as `neo4j` is this repository's shim, so the script's own `Neo4jQueryExecutor(...)` lands on the GPU
    executor (or on whatever `executor_factory` opens - the tests' CPU backend);
  * the module `common.openai_generic_assistant` is the recorded-response stub, fed per assistant name with
    the answers to replay - including a malformed first answer where a retry loop is to be exercised.
The script itself - the reference's, when its tree is present, or the mirror's dropin/test_all.py - is
imported from `tree` and its main() called unmodified; its stdout is returned.

`incident_snapshot` builds the graph such a run needs: a generated cluster in which one incident's EVENT
carries a given message (the reference's driver looks its hard-wired message up in the state graph).
"""
from __future__ import annotations

import contextlib
import importlib
import importlib.util
import io
import os
import sys

from . import dropin, executor, ingest, synth

META_URI = "bolt://10.1.0.176:7687"            # reference test_all.py:21
STATE_URI = "bolt://10.1.0.174:7687"           # reference test_all.py:22
_STAGE_PACKAGES = ("common", "find_metapath", "generate_query", "check_state", "test_all", "test_with_file")


def incident_snapshot(message: str, n_objects: int = 300, n_events: int = 30, seed: int = 7,
                      chain=("spec_volumes_persistentVolumeClaim_claimName", "spec_claimRef_uid", "spec_nfs_path")):
    """A generated cluster, re-loaded through the dump format, in which the EVENT of one incident - one whose
    Pod reaches an nfs export over claim and volume - carries `message`.  -> (snapshot, EVENT node id in it)."""
    base = synth.generate(n_objects, n_events, seed)
    lines = [__import__("json").loads(ln) for ln in ingest.dump_snapshot(base)]
    nodes = {r["id"]: r for r in lines if r["type"] == "node"}
    out, inc = {}, {}
    for r in lines:
        if r["type"] == "relationship":
            out.setdefault(r["start"]["id"], []).append(r)
            inc.setdefault(r["end"]["id"], []).append(r)

    def key(r):
        return r["properties"].get("key")
    picked = None
    for nid, n in nodes.items():
        if n["labels"] != ["Event"]:
            continue
        pods = [r["end"]["id"] for r in out.get(nid, []) if key(r) == "involvedObject_uid" and nodes[r["end"]["id"]]["labels"] == ["Pod"]]
        events = [r["end"]["id"] for r in out.get(nid, []) if r["label"] == "HasEvent"]
        for pod in pods:
            for c in (r["end"]["id"] for r in out.get(pod, []) if key(r) == chain[0]):
                for pv in (r["start"]["id"] for r in inc.get(c, []) if key(r) == chain[1]):
                    if any(key(r) == chain[2] for r in out.get(pv, [])) and events:
                        picked = events[0]
        if picked:
            break
    if picked is None:
        raise ValueError("the generated graph has no incident on a Pod with an nfs-backed claim; use more objects")
    nodes[picked]["properties"]["message"] = message
    snap = ingest.load_dump(lines)
    ev = [v for v in range(snap.n_nodes) if snap.props.node_props(v).get("message") == message]
    return snap, ev[0]


def _purge():
    for name in list(sys.modules):
        if name.split(".")[0] in _STAGE_PACKAGES:
            del sys.modules[name]


def run_driver(tree: str, snapshot, recorded: dict, script: str = "test_all.py", main_kwargs=None, executor_factory=None,
               name: str = "l3", workdir: str | None = None, pauses: bool = False):
    """Import `script` from `tree` and run its main() against `snapshot` with recorded assistant answers.
    recorded: {assistant name: {key: answer or [answers...]}} (dropin/common/openai_generic_assistant.py).
    workdir: directory to run in (test_with_file.py reads ./data/... and appends to ./output/...: with_file_workdir
    prepares one).  pauses=False turns the script's own time.sleep calls (test_with_file.py:208 waits ten seconds
    between messages) into no-ops for the run.
    Returns (stdout text, main()'s return value)."""
    executor.register_snapshot(name, snapshot)
    executor.register_uri_alias(META_URI, f"snapshot://{name}#metagraph")
    executor.register_uri_alias(STATE_URI, f"snapshot://{name}")
    dropin.install(shim_only=True, force_shim=True)
    import neo4j
    neo4j.set_executor_factory(executor_factory)
    _purge()
    mirror = dropin.HERE
    sys.path.insert(0, tree)
    try:
        stub_name = "common.openai_generic_assistant"
        if os.path.realpath(tree) != os.path.realpath(mirror):
            # a foreign tree (the reference): its own `common` package, with the assistant module replaced by the stub
            importlib.import_module("common")
            spec = importlib.util.spec_from_file_location(stub_name, os.path.join(mirror, "common", "openai_generic_assistant.py"))
            stub = importlib.util.module_from_spec(spec)
            sys.modules[stub_name] = stub
            spec.loader.exec_module(stub)
            setattr(sys.modules["common"], "openai_generic_assistant", stub)
        else:
            stub = importlib.import_module(stub_name)
        for who, table in recorded.items():
            stub.install_recorded(who, table)
        spec = importlib.util.spec_from_file_location(os.path.splitext(script)[0], os.path.join(tree, script))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[spec.name] = mod
        sink = io.StringIO()
        import time as _time
        real_sleep, here = _time.sleep, os.getcwd()
        try:
            if not pauses:
                _time.sleep = lambda *_a, **_k: None
            if workdir:
                os.chdir(workdir)
            with contextlib.redirect_stdout(sink):
                spec.loader.exec_module(mod)
                ret = mod.main(**(main_kwargs or {}))
        finally:
            _time.sleep = real_sleep
            os.chdir(here)
        return sink.getvalue(), ret
    finally:
        sys.path.remove(tree)
        neo4j.set_executor_factory(None)
        executor.unregister_uri_alias(META_URI)
        executor.unregister_uri_alias(STATE_URI)
        _purge()


def with_file_workdir(root: str, messages, input_name: str = "mixed-example-10-3.csv"):
    """The directory layout reference test_with_file.py expects (:42-43): ./data/<csv with a header row and one message
    per row> and ./output/.  Returns the path of the result file the script appends to."""
    import csv
    os.makedirs(os.path.join(root, "data"), exist_ok=True)
    os.makedirs(os.path.join(root, "output"), exist_ok=True)
    with open(os.path.join(root, "data", input_name), "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["message"])
        for m in messages:
            w.writerow([m])
    return os.path.join(root, "output", "mixed-example-10-3c-result.json")


# Progress lines both drivers print, in the order they appear: the shape of a transcript
MARKERS = ("create executor and init connection", "test find_srcKind()", "test find destKind and relevantResources",
           "JSON Error occurred", "test find_metapath()", "attempt = ", "Cypher Syntax Error occurred",
           "An unexpected error occurred", "manually generate cypher query", "there is not a STATE", "close connection")


def transcript_shape(text: str):
    """The sequence of marker lines in a driver's stdout, and of the entities whose state it checked."""
    import re
    shape = []
    for line in text.splitlines():
        if re.fullmatch(r"\w+ [0-9a-f]{8}-[0-9a-f-]{27}", line.strip()):
            shape.append("entity " + line.strip())
            continue
        for m in MARKERS:
            if m in line:
                shape.append(m if m != "attempt = " else line.strip())
                break
    return shape

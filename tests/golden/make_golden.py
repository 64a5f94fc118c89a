#!/usr/bin/env python
"""Generates tests/golden/incidents_1k.json.

Runs the REFERENCE's own stage functions, imported unmodified from /root/reference
(find_metapath/find_srckind_metapath_neo4j.py, generate_query/generate_query.py,
check_state/analyze_root_cause.py), over the CPU oracle on the seeded 1k-object graph of BASELINE
config 1, and records what they return.  The `neo4j` module they import is this repository's type
shim (the real driver is not installed); the executor they are handed is the oracle-backed one from
tests/oracle_backend.py; the assistant is the recorded-response stub.

Run in this container only (needs /root/reference):   python tests/golden/make_golden.py
The fixture is what `-m "not gpu"` tests hold the oracle path to and `-m gpu` tests hold the CUDA
path to; /root/reference itself is never read at test time.
"""
import contextlib
import io
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

# first match wins (a claim waiting for an nfs export is about the nfs volume, not the claim)
DESTS = {"nfs export": "nfs", "configmap": "ConfigMap", "secret": "Secret",
         "persistentvolumeclaim": "PersistentVolumeClaim", "nodes are available": "Node"}
INTER = {"nfs": ["PersistentVolumeClaim", "PersistentVolume"]}


def load_stage_functions(tree_root):
    """Import the four stage modules from `tree_root` (the reference or the mirror)."""
    for name in list(sys.modules):
        if name.split(".")[0] in ("common", "find_metapath", "generate_query", "check_state"):
            del sys.modules[name]
    sys.path.insert(0, tree_root)
    try:
        import find_metapath.find_srckind_metapath_neo4j as fm
        import generate_query.generate_query as gq
        import check_state.analyze_root_cause as cs
    finally:
        sys.path.remove(tree_root)
    return fm, gq, cs


def run_incidents(fm, gq, cs, meta_ex, state_ex, snap, assistant_factory, n_incidents=24):
    from oracle_backend import record_key
    out = {"kinds": None, "incidents": []}
    sink = io.StringIO()
    with contextlib.redirect_stdout(sink):
        native, external = fm.find_native_external_kinds(meta_ex)
    out["kinds"] = {"native": native, "external": external}
    msgs = snap.props.msgs
    step = max(1, len(msgs) // n_incidents)
    for i in range(0, len(msgs), step):
        msg = msgs[i]
        dest = next(v for k, v in DESTS.items() if k in msg)
        inc = {"event": i, "message": msg, "dest": dest}
        with contextlib.redirect_stdout(sink):
            inc["srcKind"] = fm.find_srcKind(state_ex, msg)
            mps = fm.find_metapath(meta_ex, inc["srcKind"], dest, INTER.get(dest, []))
        inc["metapaths"] = sorted([list(map(list, p.key())) for p in mps])
        inc["instances"] = []
        for mp in sorted(mps, key=lambda p: p.key()):
            with contextlib.redirect_stdout(sink):
                text = gq.extend_metapath_construct_string(mp)
                query = gq.human_generate_cypher_query(text, msg)
                recs = gq.run_and_filter_query(state_ex, query)
            rows = sorted(json.loads(json.dumps(record_key(r))) for r in recs)
            clues = []
            for rec in sorted(recs, key=record_key):
                with contextlib.redirect_stdout(sink):
                    report, path_clues = cs.check_statepath(state_ex, assistant_factory(), rec)
                clues.append({k: ["missing" if "there is not a STATE" in c else c.split(":")[0] for c in v]
                              for k, v in path_clues.items()})
            inc["instances"].append({"steps": [s.strip() for s in text.split(";") if s.strip()],
                                     "rows": rows, "clues": clues})
        out["incidents"].append(inc)
    return out


def main():
    ref = "/root/reference"
    if not os.path.isdir(ref):
        raise SystemExit("needs /root/reference (run in the build container)")
    from k8s_llm_rca_b200 import dropin, synth
    from oracle_backend import oracle_executor
    dropin.install(shim_only=True, force_shim=True)          # the reference does `import neo4j`
    # the reference's check_state imports its own assistant wrapper, which imports `openai`; it is never
    # instantiated here - the stub below is what gets passed in
    sys.path.insert(0, os.path.join(ROOT, "k8s-llm-rca_b200", "dropin"))
    from common.openai_generic_assistant import OpenAIGenericAssistant as Stub
    sys.path.pop(0)
    fm, gq, cs = load_stage_functions(ref)
    assert fm.__file__.startswith(ref) and gq.__file__.startswith(ref) and cs.__file__.startswith(ref)
    snap = synth.generate(1000, 100, synth.config_seed(0))
    out = run_incidents(fm, gq, cs, oracle_executor(snap.derive_metagraph()), oracle_executor(snap), snap,
                        lambda: Stub({"summarize the root cause": "{}"}))
    out["generator"] = {"objects": 1000, "events": 100, "seed": synth.config_seed(0),
                        "source": "reference stage functions, unmodified, over oracle/csr_oracle.c"}
    path = os.path.join(HERE, "incidents_1k.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    n_rows = sum(len(x["rows"]) for inc in out["incidents"] for x in inc["instances"])
    print(f"wrote {path}: {len(out['incidents'])} incidents, {n_rows} instance rows")


if __name__ == "__main__":
    main()

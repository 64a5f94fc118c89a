"""BASELINE config 4 at one GPU's share: 10 M-object graph, <= 5 hops, seeds in chunks.
Counts for every chunk (count_only), rows for the first chunk.  Prints one JSON line per chunk."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from k8s_llm_rca_b200 import engine as E, synth  # noqa: E402

n_obj = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
chunk = int(sys.argv[2]) if len(sys.argv) > 2 else 250_000
n_chunks = int(sys.argv[3]) if len(sys.argv) > 3 else 4
t0 = time.time()
snap = synth.generate(n_obj, n_obj // 100, synth.config_seed(3))
print(json.dumps({"generated_s": round(time.time() - t0, 1), "nodes": snap.n_nodes, "rels": snap.n_rels,
                  "device_MB": round(snap.device_bytes() / 1e6)}), flush=True)
eng = E.Engine(snap)
K = snap.kind_index
pod, node = K("Pod"), K("Node")
excl = [K("Event"), K("Namespace")]
pods = snap.nodes_of_kind(pod)
for hops in (4, 5):
    plan = E.varlen_plan("und", hops, dest_kind=node, excl_kinds=excl, entity_only=True, pos_kind=[pod], count_only=True)
    for c in range(n_chunks):
        seeds = pods[c * chunk:(c + 1) * chunk]
        for rep in range(2):
            r = eng.expand(plan, seeds)
        alg = sum(12 * r.frontier[s - 1] + 4 * r.scanned[s] for s in range(1, hops + 1))
        print(json.dumps({"hops": hops, "chunk": c, "seeds": len(seeds), "F": r.frontier, "E": r.scanned, "N": r.emitted,
                          "kernel_ms": round(r.kernel_ms, 3), "level_ms": [round(x, 3) for x in r.level_ms],
                          "instances_per_s": round(r.total / (r.kernel_ms * 1e-3)), "launches": r.launches,
                          "retries": r.retries, "count_pass_GBps": round(alg / (r.kernel_ms * 1e-3) / 1e9, 1)}), flush=True)
plan = E.varlen_plan("und", 4, dest_kind=node, excl_kinds=excl, entity_only=True, pos_kind=[pod])
for rep in range(2):
    r = eng.expand(plan, pods[:chunk])
print(json.dumps({"hops": 4, "with_rows": True, "seeds": chunk, "N": r.emitted, "kernel_ms": round(r.kernel_ms, 3),
                  "instances_per_s": round(r.total / (r.kernel_ms * 1e-3)), "retries": r.retries}), flush=True)

"""Profiling driver: the enumeration plan of the config-3 workload, a few calls, counters printed.
Run plain first, then under ncu (see profiles/README.md)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from k8s_llm_rca_b200 import engine as E  # noqa: E402
from k8s_llm_rca_b200 import workloads  # noqa: E402

n_obj = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
chains = len(sys.argv) > 3 and sys.argv[3] == "chains"
w = workloads.Config3(n_objects=n_obj, n_events=int(n_obj * 0.108696))
eng = E.Engine(w.snap)
eng.reserve(48_000_000, 160_000_000)   # no overflow retries: call k, level s is kernel 4k+s-1
pods, ev, ts = w.batch(0)
for i in range(reps):
    r = eng.expand(w.varlen, pods, ts)
    if chains:
        for name, (_, _, plan) in w.chains.items():
            rc = eng.expand(plan, ev, ts)
            print(json.dumps({"plan": name, "F": rc.frontier, "E": rc.scanned, "N": rc.emitted,
                              "level_ms": [round(x, 4) for x in rc.level_ms], "kernel_ms": round(rc.kernel_ms, 4)}))
print(json.dumps({"plan": "varlen", "F": r.frontier, "E": r.scanned, "N": r.emitted, "row_len": r.row_len,
                  "level_ms": [round(x, 4) for x in r.level_ms], "kernel_ms": round(r.kernel_ms, 4),
                  "level_bytes": [0] + [r.level_bytes(s) for s in range(1, r.levels + 1)],
                  "launches": r.launches, "retries": r.retries}))

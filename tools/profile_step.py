"""Profiling driver: whole config-3 steps as bench.py submits them (one rca_expand_many of the enumeration
and the five instance queries).  Run plain first, then under ncu (see profiles/README.md)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from k8s_llm_rca_b200 import engine as E  # noqa: E402
from k8s_llm_rca_b200 import workloads  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
w = workloads.Config3()
eng = E.Engine(w.snap)
eng.reserve(48_000_000, 160_000_000)
pods, ev, ts = w.batch(0)
for i in range(reps):
    tot, res = w.run(eng, pods, ev, ts)
print("instances", tot, [(n, r.launches) for n, r in res])

"""Where the time of a batched incident replay goes (10 k incidents of the config-3 graph): three timed runs, then a
cProfile of one.  python tools/latency_batch_profile.py"""
import cProfile
import os
import pstats
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from k8s_llm_rca_b200 import executor, replay, workloads
w = workloads.Config3()
executor.register_snapshot("bench", w.snap)
sx = executor.GpuQueryExecutor("snapshot://bench"); mx = executor.GpuQueryExecutor("snapshot://bench#metagraph")
rp = replay.IncidentReplayer(sx, mx)
rng = np.random.default_rng(5); msgs = w.snap.props.msgs
draw = [msgs[i] for i in rng.integers(0, len(msgs), size=10000)]
rp.run_batch_arrays(draw)
for _ in range(3):
    t0 = time.perf_counter(); out = rp.run_batch_arrays(draw); print("batch ms", (time.perf_counter() - t0) * 1e3, out.n_rows, np.percentile(out.done_ms[out.done_ms>0], [50, 99]))
pr = cProfile.Profile(); pr.enable(); rp.run_batch_arrays(draw); pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)

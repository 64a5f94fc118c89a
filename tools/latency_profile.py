"""Where the per-incident latency goes: cProfile of IncidentReplayer.run_one over the config-3 graph
(or a smaller one: python tools/latency_profile.py [n_objects])."""
import cProfile
import os
import pstats
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from k8s_llm_rca_b200 import executor, replay, synth

n_obj = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
snap = synth.generate(n_obj, int(n_obj * 0.108696), synth.config_seed(2))
executor.register_snapshot("lat", snap)
sx = executor.GpuQueryExecutor("snapshot://lat")
mx = executor.GpuQueryExecutor("snapshot://lat#metagraph")
rp = replay.IncidentReplayer(sx, mx)
msgs = snap.props.msgs
rng = np.random.default_rng(5)
draw = [msgs[i] for i in rng.integers(0, len(msgs), size=300)]
for m in draw[:20]:
    rp.run_one(m)
lat = []
for m in draw[20:220]:
    t0 = time.perf_counter()
    rp.run_one(m)
    lat.append((time.perf_counter() - t0) * 1e3)
print("p50 %.3f ms  p99 %.3f ms" % (np.percentile(lat, 50), np.percentile(lat, 99)))
pr = cProfile.Profile()
pr.enable()
for m in draw[220:300]:
    rp.run_one(m)
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("cumtime").print_stats(45)

"""Host-side cost of one rca_expand_many submission: the config-3 plan set over ONE seed each (GPU work is
negligible), wall time per call.  python tools/host_overhead.py"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from k8s_llm_rca_b200 import engine as E, synth, workloads

snap = synth.generate(20000, 2174, synth.config_seed(2))
w = workloads.Config3(snap=snap)
eng = E.Engine(snap)
pods, ev, ts = w.batch(0)
for n in (1, len(pods)):
    plans = [w.varlen] + [p for (_, _, p) in w.chains.values()]
    seeds = [pods[:n]] + [ev[:n]] * len(w.chains)
    tss = [ts[:n]] * len(plans)
    for i in range(20):
        eng.expand_many(plans, seeds, tss)
    t0 = time.perf_counter()
    for i in range(200):
        eng.expand_many(plans, seeds, tss)
    t1 = time.perf_counter()
    r = eng.expand_many(plans, seeds, tss)
    print(f"seeds={n}: {1e6 * (t1 - t0) / 200:.1f} us per submission of {len(plans)} plans; device ms per plan {[round(x.kernel_ms, 4) for x in r]}")

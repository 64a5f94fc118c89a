"""Config-3 instance queries (typed chains) with the unit-step tables on and off: device time per chain,
L2 flushed before every run.  python tools/chain_ab.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from k8s_llm_rca_b200 import engine as E, workloads

w = workloads.Config3()
eng = E.Engine(w.snap)
pods, ev, ts = w.batch(0)
eng.reserve(48_000_000, 160_000_000)
flush = torch.empty(1 << 28, dtype=torch.uint8, device="cuda")
for on in (True, False, True, False):
    eng.set_unit_steps(on)
    out = {}
    for name, (_, _, plan) in w.chains.items():
        t = []
        for i in range(12):
            flush.zero_()
            torch.cuda.synchronize()
            r = eng.expand(plan, ev, ts)
            if i >= 3:
                t.append(r.kernel_ms)
        out[name] = (round(float(np.median(t)), 4), r.total, r.launches)
    print(f"unit={'on ' if on else 'off'}", out, "sum_ms", round(sum(v[0] for v in out.values()), 4))

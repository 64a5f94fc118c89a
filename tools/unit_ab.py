"""Config-3 enumeration with the unit-step kernels on and off: per-level device time (CUDA events inside
the engine), L2 flushed before every run.  python tools/unit_ab.py"""
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
    lv, tot = [], []
    for i in range(25):
        flush.zero_()
        torch.cuda.synchronize()
        r = eng.expand(w.varlen, pods, ts)
        if i >= 5:
            lv.append(r.level_ms)
            tot.append(r.kernel_ms)
    lv = np.median(np.array(lv), axis=0)
    print(f"unit={'on ' if on else 'off'} levels_ms={[round(float(x), 3) for x in lv]} kernel_ms={np.median(tot):.3f} "
          f"N={r.emitted} launches={r.launches}")

"""Where a config-3 step's time goes: the six plans together, the enumeration alone, the five instance
queries alone - device time between CUDA events around the call (as bench.py measures) and host wall time.
python tools/step_parts.py"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from k8s_llm_rca_b200 import engine as E, workloads

w = workloads.Config3()
eng = E.Engine(w.snap)
stream = torch.cuda.Stream(priority=-1)
eng.set_stream(stream.cuda_stream)
pods, ev, ts = w.batch(0)
with torch.cuda.stream(stream):
    d_pods, d_ev, d_ts = (torch.from_numpy(x).cuda() for x in (pods, ev, ts))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    chains = [c[2] for c in w.chains.values()]
    cases = {
        "all six plans": lambda: eng.expand_many([w.varlen] + chains, [d_pods] + [d_ev] * 5, [d_ts] * 6),
        "enumeration only (expand)": lambda: eng.expand(w.varlen, d_pods, d_ts),
        "enumeration only (expand_many)": lambda: eng.expand_many([w.varlen], [d_pods], [d_ts]),
        "five chains only": lambda: eng.expand_many(chains, [d_ev] * 5, [d_ts] * 5),
        "one chain (expand)": lambda: eng.expand(chains[0], d_ev, d_ts),
    }
    for name, fn in cases.items():
        for _ in range(5):
            fn()
        dev, wall = [], []
        for _ in range(40):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            t0 = time.perf_counter()
            fn()
            wall.append((time.perf_counter() - t0) * 1e3)
            e1.record(stream)
            torch.cuda.synchronize()
            dev.append(e0.elapsed_time(e1))
        print(f"{name:34s} device {np.median(dev):.3f} ms   host wall {np.median(wall):.3f} ms")
        r = fn()
        for x in (r if isinstance(r, list) else [r]):
            print("      plan: kernel_ms %.3f  level_ms %s" % (x.kernel_ms, [round(v, 3) for v in x.level_ms]))

"""Does the per-level device time settle over repetitions (clock ramp) ?"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from k8s_llm_rca_b200 import engine as E, workloads
w = workloads.Config3()
eng = E.Engine(w.snap)
pods, ev, ts = w.batch(0)
import subprocess
for i in range(40):
    r = eng.expand(w.varlen, pods, ts)
    if i % 4 == 0:
        clk = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.mem,power.draw", "--format=csv,noheader"], capture_output=True, text=True).stdout.strip()
        print(i, [round(x, 3) for x in r.level_ms], round(r.kernel_ms, 3), clk)
import torch
x = torch.empty(1 << 28, dtype=torch.uint8, device="cuda")
for i in range(5):
    x.zero_(); torch.cuda.synchronize()
    r = eng.expand(w.varlen, pods, ts)
    print("after flush", [round(x, 3) for x in r.level_ms], round(r.kernel_ms, 3))

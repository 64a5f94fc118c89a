import sys, os
sys.path.insert(0, os.getcwd())
from k8s_llm_rca_b200 import engine as E, workloads
w = workloads.Config3()
eng = E.Engine(w.snap); eng.reserve(48_000_000, 160_000_000)
pods, ev, ts = w.batch(0)
K = w.snap.kind_index
for name, plan, t in (("with states", w.varlen, ts), ("no states", E.varlen_plan("und", 4, dest_kind=K("Node"), excl_kinds=w.excl, entity_only=True, pos_kind=[K("Pod")]), None), ("count only", E.varlen_plan("und", 4, dest_kind=K("Node"), excl_kinds=w.excl, entity_only=True, pos_kind=[K("Pod")], count_only=True), None)):
    for i in range(3): r = eng.expand(plan, pods, t)
    print(name, [round(x,3) for x in r.level_ms], round(r.kernel_ms,3))

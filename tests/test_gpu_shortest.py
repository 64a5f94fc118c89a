"""rca_plan_t::shortest_only - the rule find_metapath applies to what its statement returned (reference
find_metapath/find_srckind_metapath_neo4j.py:153-154: minLen = min(len(path)); keep the paths of that length) done by
the engine: the launches of every longer length return at once.  The kept rows must be exactly the oracle's rows of the
minimal length, for every pair of kinds, directed and undirected, through both kernel families; and through the
mirror's find_metapath the executor's answer is unchanged."""
import numpy as np
import pytest

from conftest import sort_rows

pytestmark = pytest.mark.gpu
KINDS = ["Pod", "ReplicaSet", "Deployment", "Service", "PersistentVolumeClaim", "PersistentVolume", "ConfigMap", "Secret", "nfs", "Node"]


@pytest.mark.parametrize("family", ["fused", "split"])
def test_only_the_shortest_length_is_produced(snap_small, csr_oracle_lib, family):
    from k8s_llm_rca_b200 import engine
    from k8s_llm_rca_b200.synth import T_BASE, DAY
    snap = snap_small
    orc = csr_oracle_lib.CsrOracle(snap)
    eng = engine.Engine(snap)
    eng.set_fused_threshold(1 << 40 if family == "fused" else 0)
    K = snap.kind_index
    excl = [K("Event"), K("Namespace")]
    lengths = set()
    for direction in ("und", "out"):
        for src in KINDS:
            seeds = snap.nodes_of_kind(K(src)).astype(np.int32)
            ts = (T_BASE + np.random.default_rng(3).integers(0, 30 * DAY, size=len(seeds))).astype(np.int64)
            for dst in KINDS:
                kw = dict(dest_kind=K(dst), excl_kinds=excl, entity_only=True)
                want, _ = orc.varlen(seeds, direction, 4, **kw)
                res = eng.expand(engine.varlen_plan(direction, 4, shortest_only=True, with_states=True, **kw), seeds, ts)
                if not want:
                    assert res.total == 0
                    continue
                m = min(want)                                   # find_metapath: minLen = min(len(path) for path in paths)
                lengths.add(m)
                assert [s for s in range(1, 5) if res.emitted[s]] == [m], (direction, src, dst)
                kept = sort_rows(res.rows(m))                   # copied out before the engine runs the next plan
                emitted, frontier = list(res.emitted), list(res.frontier)
                assert np.array_equal(kept[:, :2 * m + 1], sort_rows(want[m]))
                full = eng.expand(engine.varlen_plan(direction, 4, with_states=True, **kw), seeds, ts)
                assert np.array_equal(sort_rows(full.rows(m)), kept)                       # state columns too
                assert frontier[m + 1:] == [0] * (4 - m)        # nothing was expanded past the shortest length
                cnt = eng.expand(engine.varlen_plan(direction, 4, shortest_only=True, count_only=True, **kw), seeds)
                assert list(cnt.emitted) == emitted
    assert lengths >= {1, 2, 3}
    eng.close()

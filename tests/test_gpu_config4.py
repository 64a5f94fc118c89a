"""BASELINE config 4's shape - undirected Pod -> Node enumeration of up to FIVE hops on a graph with hub Nodes
(200 k objects: a thousand Nodes of a hundred Pods each) - against the CPU oracle:
  * count-only plans, through the count / scan / fill kernels (where the level before the last takes the final
    unit step in its count pass and the last level is never launched) and through the single-launch kernels:
    per-length counts equal the oracle's row counts, for several thousand seeds;
  * the same plan with rows: bit-exact sorted row sets for a seed sample, both kernel families;
  * a count-only run reports what the run with rows emits."""
import numpy as np
import pytest

from conftest import sort_rows

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def shape4(csr_oracle_lib):
    from k8s_llm_rca_b200 import engine, synth
    snap = synth.generate(200_000, 2000, seed=synth.config_seed(3))
    eng = engine.Engine(snap)
    yield snap, eng, csr_oracle_lib.CsrOracle(snap)
    eng.close()


@pytest.mark.parametrize("family", ["split", "fused"])
def test_five_hops_counts_and_rows(shape4, family):
    from k8s_llm_rca_b200 import engine
    snap, eng, orc = shape4
    eng.set_fused_threshold(0 if family == "split" else 1 << 40)
    K = snap.kind_index
    pod, node = K("Pod"), K("Node")
    kw = dict(dest_kind=node, excl_kinds=[K("Event"), K("Namespace")], entity_only=True)
    pods = snap.nodes_of_kind(pod)
    seeds = np.ascontiguousarray(pods[::len(pods) // 3000][:3000])
    counted = eng.expand(engine.varlen_plan("und", 5, pos_kind=[pod], count_only=True, **kw), seeds)
    ref, _ = orc.varlen(seeds, "und", 5, **kw)
    want = [len(ref[s]) if s in ref else 0 for s in range(6)]
    assert list(counted.emitted) == want and counted.total == sum(want)
    assert want[5] > 10_000_000 and want[3] > 0              # hubs: thousands of five-hop paths per seed
    with pytest.raises(ValueError):
        counted.rows(5)                                       # a count-only run holds no rows
    # <= 4 hops, counted, on all of these seeds as well (the other configuration bench.py's sharded leg runs)
    c4 = eng.expand(engine.varlen_plan("und", 4, pos_kind=[pod], count_only=True, **kw), seeds)
    assert list(c4.emitted) == want[:5]
    # rows for a sample of the seeds
    sample = np.ascontiguousarray(seeds[::10])
    res = eng.expand(engine.varlen_plan("und", 5, pos_kind=[pod], **kw), sample)
    ref_s, _ = orc.varlen(sample, "und", 5, **kw)
    got = res.all_rows()
    assert sorted(got) == sorted(ref_s)
    for s in ref_s:
        assert np.array_equal(sort_rows(got[s]), sort_rows(ref_s[s])), s
    # without pos_kind (seed kind not declared) and without pruning: same rows
    plain = eng.expand(engine.varlen_plan("und", 5, prune=False, **kw), sample[:40])
    ref_p, _ = orc.varlen(sample[:40], "und", 5, **kw)
    got_p = plain.all_rows()
    for s in ref_p:
        assert np.array_equal(sort_rows(got_p[s]), sort_rows(ref_p[s])), s


def test_five_hops_with_states_and_sharding(shape4):
    """State columns at five hops, and linearity over contiguous shards (what the multi-GPU split relies on)."""
    from k8s_llm_rca_b200 import engine, parallel
    from k8s_llm_rca_b200.synth import T_BASE, DAY
    snap, eng, orc = shape4
    eng.set_fused_threshold(0)
    K = snap.kind_index
    pod, node = K("Pod"), K("Node")
    kw = dict(dest_kind=node, excl_kinds=[K("Event"), K("Namespace")], entity_only=True)
    pods = snap.nodes_of_kind(pod)
    seeds = np.ascontiguousarray(pods[::len(pods) // 240][:240])
    ts = (T_BASE + np.random.default_rng(4).integers(0, 30 * DAY, size=len(seeds))).astype(np.int64)
    plan = engine.varlen_plan("und", 5, pos_kind=[pod], with_states=True, **kw)
    whole = {s: r.copy() for s, r in eng.expand(plan, seeds, ts).all_rows().items()}
    want, _, _ = orc.varlen_step(seeds, ts, "und", 5, **kw)
    for s in want:
        assert np.array_equal(sort_rows(whole[s]), sort_rows(want[s])), s
    parts = {s: [] for s in whole}
    for r in range(3):
        sd, st, _ = parallel.shard_seeds(snap, seeds, r, 3, "und", True, seed_ts=ts)
        for s, rows in eng.expand(plan, sd, st).all_rows().items():
            parts[s].append(rows.copy())
    for s in whole:                                           # the shards' rows together are the single run's rows
        assert np.array_equal(sort_rows(np.concatenate(parts[s])), sort_rows(whole[s])), s

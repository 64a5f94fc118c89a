"""Parity at BASELINE.json's full size (config 3: 1 M objects, 100 k seed Pods), through properties
that do not need the oracle to enumerate everything:
  * exact agreement with the C oracle on a random sample of the seeds;
  * linearity: the rows of the whole batch are the disjoint union of the rows of its shards;
  * pruning is result-neutral: counts with and without schema pruning are equal;
  * determinism: two runs give byte-identical buffers;
  * symmetry: every undirected Pod -> Node path reversed is an undirected Node -> Pod path.
"""
import numpy as np
import pytest

from conftest import rows_to_set, sort_rows

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def big():
    from k8s_llm_rca_b200 import engine, workloads
    w = workloads.Config3()
    eng = engine.Engine(w.snap)
    yield w, eng
    eng.close()


def test_sample_of_seeds_matches_oracle(big, csr_oracle_lib):
    w, eng = big
    pods, ev, ts = w.batch(0)
    orc = csr_oracle_lib.CsrOracle(w.snap)
    pick = np.sort(np.random.default_rng(12).choice(len(pods), size=3000, replace=False))
    res = eng.expand(w.varlen, pods[pick], ts[pick])
    ref, _ = orc.varlen(pods[pick], "und", w.max_hops, dest_kind=w.snap.kind_index("Node"), excl_kinds=w.excl,
                        entity_only=True)
    all_rows = res.all_rows()                       # copy out before the engine's arena is reused
    assert rows_to_set(all_rows) == rows_to_set(ref)
    for name, (ets, dirs, plan) in w.chains.items():
        got = eng.expand(plan, ev[pick], ts[pick]).rows(len(ets))
        want, _ = orc.chain(ev[pick], ets, [{"out": 0, "in": 1}[d] for d in dirs])
        assert np.array_equal(sort_rows(got[:, :2 * len(ets) + 1]), sort_rows(want)), name
    # fused state columns against the oracle's interval test
    s = 3
    got = sort_rows(all_rows[s])
    t = ts[pick][np.searchsorted(pods[pick], got[:, 0])]
    miss = np.zeros(len(got), np.int64)
    for d in range(s + 1):
        first, cnt, _ = orc.states(got[:, d], t)
        assert np.array_equal(got[:, 2 * s + 1 + d], first)
        miss |= (cnt == 0).astype(np.int64) << d
    assert np.array_equal(got[:, 3 * s + 2].astype(np.int64), miss)


def test_full_batch_linearity_pruning_determinism(big):
    from k8s_llm_rca_b200 import engine
    w, eng = big
    pods, ev, ts = w.batch(0)
    full = eng.expand(w.varlen, pods, ts)
    emitted = list(full.emitted)
    rows3 = full.rows(3).copy()
    again = eng.expand(w.varlen, pods, ts)
    assert list(again.emitted) == emitted and np.array_equal(again.rows(3), rows3)          # determinism
    # shards: contiguous quarters of the seed list; their rows concatenate to the full result, in order
    parts, tot = [], np.zeros(len(emitted), np.int64)
    for lo in range(0, len(pods), len(pods) // 4):
        r = eng.expand(w.varlen, pods[lo:lo + len(pods) // 4], ts[lo:lo + len(pods) // 4])
        tot += np.array(r.emitted)
        parts.append(r.rows(3).copy())
    assert tot.tolist() == emitted
    assert np.array_equal(sort_rows(np.concatenate(parts)), sort_rows(rows3))
    # schema pruning off: same counts, more candidates scanned
    K = w.snap.kind_index
    plain = engine.varlen_plan("und", w.max_hops, dest_kind=K("Node"), excl_kinds=w.excl, entity_only=True,
                               prune=False, with_states=True)
    slow = eng.expand(plain, pods[:20000], ts[:20000])
    slow3 = slow.rows(3).copy()
    fast = eng.expand(w.varlen, pods[:20000], ts[:20000])
    assert list(slow.emitted) == list(fast.emitted) and sum(slow.scanned) > 3 * sum(fast.scanned)
    assert np.array_equal(sort_rows(slow3), sort_rows(fast.rows(3)))


def test_undirected_paths_are_symmetric(big):
    from k8s_llm_rca_b200 import engine
    w, eng = big
    K = w.snap.kind_index
    pod, node = K("Pod"), K("Node")
    nodes = w.snap.nodes_of_kind(node)[:40]
    back = engine.varlen_plan("und", 3, dest_kind=pod, excl_kinds=w.excl, entity_only=True, pos_kind=[node])
    rb = eng.expand(back, nodes).all_rows()
    # the Pods those paths end at, expanded forwards, must contain exactly the reversed paths
    ends = np.unique(np.concatenate([rows[:, s] for s, rows in rb.items()]))
    fwd = engine.varlen_plan("und", 3, dest_kind=node, excl_kinds=w.excl, entity_only=True, pos_kind=[pod])
    rf = eng.expand(fwd, ends).all_rows()
    want = set()
    for s, rows in rb.items():
        for r in rows.tolist():
            want.add((tuple(r[:s + 1][::-1]), tuple(r[s + 1:2 * s + 1][::-1])))
    got = {p for p in rows_to_set(rf) if p[0][-1] in set(nodes.tolist())}
    assert got == want and len(want) > 1000

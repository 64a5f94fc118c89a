"""Randomised parity: small random multigraphs - parallel relationships, self-loops, kinds with no nodes, edge types
that are functional on one side and not on the other, HasState intervals of every count - and random plans
(direction, 1..8 hops, destination or none, excluded and intermediate kinds, node-simple or relationship-unique
patterns, a declared seed kind, state columns, shortest-only, count-only) through every kernel family, against the C
oracle.  The generated cluster graphs of the other tests share one schema; these graphs have none."""
import numpy as np
import pytest

from conftest import sort_rows

pytestmark = pytest.mark.gpu
KINDS = ["A", "B", "C", "D", "Event", "Namespace"]
T0, H = 1_700_000_000_000, 3_600_000


def random_graph(rng):
    from k8s_llm_rca_b200.props import TableProps
    from k8s_llm_rca_b200.snapshot import Kind, Snapshot
    counts = [int(rng.integers(0, 7)) for _ in KINDS]
    counts[0] = max(counts[0], 2)
    kinds = [Kind(k, k, "NativeEntity") for k in KINDS] + [Kind(k, k.upper() + "_S", "NativeState", True) for k in KINDS[:4]]
    n_state = [int(rng.integers(0, 3 * max(1, counts[i]))) for i in range(4)]
    all_counts = counts + n_state
    start = np.concatenate([[0], np.cumsum(all_counts)])
    ets, idx, es, ed, ee, tmin, tmax = [], {}, [], [], [], [], []

    def add(rel, sk, dk, key, s, d, lo=0, hi=0):
        k = (rel, sk, dk, key)
        if k not in idx:
            idx[k] = len(ets)
            ets.append(k)
        es.append(int(start[sk] + s)); ed.append(int(start[dk] + d)); ee.append(idx[k]); tmin.append(lo); tmax.append(hi)
    n_edges = int(rng.integers(3, 40))
    for _ in range(n_edges):
        sk, dk = int(rng.integers(0, 6)), int(rng.integers(0, 6))
        if counts[sk] == 0 or counts[dk] == 0:
            continue
        t = int(rng.integers(0, 3))
        s, d = int(rng.integers(0, counts[sk])), int(rng.integers(0, counts[dk]))
        add("R%d" % t, sk, dk, "k%d" % t, s, d)
        if rng.random() < 0.2:
            add("R%d" % t, sk, dk, "k%d" % t, s, d)                 # a parallel relationship
    for i in range(4):                                              # functional edge types: each A-node has at most one
        for s in range(counts[i]):
            j = (i + 1) % 4
            if counts[j] and rng.random() < 0.8:
                add("F", i, j, "f", s, int(rng.integers(0, counts[j])))
    for i in range(4):
        for k in range(n_state[i]):
            if counts[i] == 0:
                continue
            lo = int(rng.integers(0, 20))
            add("HasState", i, 6 + i, None, int(rng.integers(0, counts[i])), k, T0 + lo * H, T0 + (lo + int(rng.integers(1, 8))) * H)
    snap = Snapshot.build(kinds, all_counts, ets, es, ed, ee, tmin, tmax)
    snap.props = TableProps(snap, [{"kind": kinds[int(snap.kind_of(v))].name} for v in range(snap.n_nodes)])
    return snap, counts


def test_random_graphs_and_plans_against_the_oracle(csr_oracle_lib):
    from k8s_llm_rca_b200 import engine
    rng = np.random.default_rng(20240)
    compared = rows_seen = 0
    for g in range(200):
        snap, counts = random_graph(rng)
        if snap.n_rels == 0:
            continue
        orc = csr_oracle_lib.CsrOracle(snap)
        eng = engine.Engine(snap)
        for fam in (0, 1 << 40):
            eng.set_fused_threshold(fam)
            unit_on, wide_on = bool(rng.random() < 0.8), bool(rng.random() < 0.5)
            eng.set_unit_steps(unit_on)                         # the result-neutral switches, in every combination
            eng.set_wide_entries(wide_on)
            for _ in range(6):
                direction = ["out", "in", "und"][int(rng.integers(0, 3))]
                hops = int(rng.integers(1, 9))
                src = int(rng.choice([k for k in range(4) if counts[k]]))
                dest = int(rng.integers(-1, 4))
                node_simple = bool(rng.random() < 0.8)
                kw = dict(dest_kind=dest, excl_kinds=[4, 5] if rng.random() < 0.7 else [], entity_only=bool(rng.random() < 0.7),
                          node_simple=node_simple, min_hops=int(rng.integers(1, hops + 1)))
                if rng.random() < 0.3:
                    kw["inter_kinds"] = [int(x) for x in rng.choice(4, size=int(rng.integers(1, 3)), replace=False)]
                if rng.random() < 0.5:
                    kw["pos_kind"] = [src] + [-1] * hops
                    if rng.random() < 0.3:                      # a kind required somewhere inside the path as well
                        kw["pos_kind"][int(rng.integers(1, hops + 1))] = int(rng.integers(0, 4))
                seeds = snap.nodes_of_kind(src).astype(np.int32)
                if rng.random() < 0.3:
                    seeds = np.concatenate([seeds, seeds[:1], [-1, snap.n_nodes + 3]]).astype(np.int32)     # repeated and invalid seeds
                ok = seeds[(seeds >= 0) & (seeds < snap.n_nodes)]
                ts = (T0 + rng.integers(-2, 30, size=len(seeds)) * H).astype(np.int64)
                want, _, _ = orc.varlen_step(ok, ts[(seeds >= 0) & (seeds < snap.n_nodes)], direction, hops, **kw)
                for mode in (False, True):
                    prune = bool(rng.random() < 0.7)
                    try:
                        res = eng.expand(engine.varlen_plan(direction, hops, with_states=mode, prune=prune, **kw), seeds, ts)
                    except engine.RcaError as ex:
                        raise AssertionError(f"{ex}: graph {g}, {direction} {hops} hops, {kw}, states {mode}, prune {prune}, "
                                             f"fused {fam != 0}, unit {unit_on}, wide {wide_on}, {len(seeds)} seeds") from None
                    got = res.all_rows()
                    assert sorted(got) == sorted(want), (g, direction, hops, kw)
                    for s in want:
                        w = want[s] if mode else want[s][:, :2 * s + 1]
                        assert np.array_equal(sort_rows(got[s]), sort_rows(w)), (g, direction, hops, kw, s, mode)
                        rows_seen += len(w)
                short = eng.expand(engine.varlen_plan(direction, hops, shortest_only=True, count_only=bool(rng.random() < 0.5), **kw), seeds)
                m = min(want) if want else None
                assert [s for s in range(1, hops + 1) if short.emitted[s]] == ([m] if m else []), (g, direction, hops, kw)
                if m:
                    assert short.emitted[m] == len(want[m])
                compared += 1
        eng.close()
    assert compared > 1600 and rows_seen > 16000


def test_random_typed_chains_against_the_oracle(csr_oracle_lib):
    """Random typed chains (the instance query's shape: one edge type and one direction per step, reversed steps
    included) over the same random graphs, one by one - as one depth-first launch and level by level - and several at
    once in one submission (chain_batch), with full state columns and with the mask alone."""
    from k8s_llm_rca_b200 import engine
    rng = np.random.default_rng(777)
    compared = rows_seen = 0
    for g in range(120):
        snap, counts = random_graph(rng)
        if snap.n_rels == 0:
            continue
        orc = csr_oracle_lib.CsrOracle(snap)
        eng = engine.Engine(snap)
        chains = []
        for _ in range(8):
            cur = int(rng.choice([k for k in range(4) if counts[k]]))
            k0, ets, dirs = cur, [], []
            for _step in range(int(rng.integers(1, 6))):
                opts = [(t, "out", e.dst_kind) for t, e in enumerate(snap.etypes) if e.src_kind == cur] + \
                       [(t, "in", e.src_kind) for t, e in enumerate(snap.etypes) if e.dst_kind == cur]
                if not opts:
                    break
                t, d, cur = opts[int(rng.integers(0, len(opts)))]
                ets.append(t); dirs.append(d)
            if ets:
                chains.append((k0, ets, dirs))
        for walk in (64, 0):                                    # one launch per chain / one launch per step
            eng.set_chain_walk_fanout(walk)
            for k0, ets, dirs in chains:
                seeds = snap.nodes_of_kind(k0).astype(np.int32)
                ts = (T0 + rng.integers(-2, 30, size=len(seeds)) * H).astype(np.int64)
                L = len(ets)
                want, _ = orc.chain_step(seeds, ts, ets, [1 if d == "in" else 0 for d in dirs])
                full = eng.expand(engine.chain_plan(ets, dirs, with_states=True), seeds, ts)
                assert full.total == len(want)
                assert np.array_equal(sort_rows(full.rows(L)), sort_rows(want)), (g, ets, dirs, walk)
                mask = eng.expand(engine.chain_plan(ets, dirs, with_states="mask"), seeds, ts).rows(L)
                w = sort_rows(want)
                assert np.array_equal(sort_rows(mask), np.concatenate([w[:, :2 * L + 1], w[:, 3 * L + 2:]], axis=1) if len(w) else mask)
                plain = eng.expand(engine.chain_plan(ets, dirs), seeds)
                assert np.array_equal(sort_rows(plain.rows(L)), w[:, :2 * L + 1] if len(w) else plain.rows(L))
                assert eng.expand(engine.chain_plan(ets, dirs, count_only=True), seeds).total == len(want)
                compared += 1
                rows_seen += len(want)
        # all chains of one source kind in one submission
        eng.set_chain_walk_fanout(64)
        eng.set_graphs(bool(rng.random() < 0.5))
        eng.set_unit_steps(bool(rng.random() < 0.8))
        by_kind = {}
        for c in chains:
            by_kind.setdefault(c[0], []).append(c)
        for k0, group in by_kind.items():
            seeds = snap.nodes_of_kind(k0).astype(np.int32)
            ts = (T0 + rng.integers(0, 25, size=len(seeds)) * H).astype(np.int64)
            plans = [engine.chain_plan(ets, dirs, with_states=True) for _, ets, dirs in group]
            out = eng.expand_many(plans, [seeds] * len(plans), [ts] * len(plans))
            rows = [r.rows(len(group[i][1])).copy() for i, r in enumerate(out)]
            for (_, ets, dirs), got in zip(group, rows):
                want, _ = orc.chain_step(seeds, ts, ets, [1 if d == "in" else 0 for d in dirs])
                assert np.array_equal(sort_rows(got), sort_rows(want)), (g, ets, dirs, "batched")
        eng.close()
    assert compared > 1000 and rows_seen > 3000

"""Randomised parity: small random multigraphs - parallel relationships, self-loops, kinds with no nodes, edge types
that are functional on one side and not on the other, HasState intervals of every count - and random plans
(direction, 1..8 hops, destination or none, excluded and intermediate kinds, node-simple or relationship-unique
patterns, a declared seed kind, state columns, shortest-only, count-only) through every kernel family, against the C
oracle.  The generated cluster graphs of the other tests share one schema; these graphs have none."""
import numpy as np
import pytest

from conftest import sort_rows

import os

pytestmark = pytest.mark.gpu
SEED = int(os.environ.get("RCA_FUZZ_SEED", "0"))      # other random streams: RCA_FUZZ_SEED=1, 2, ... (a hunt, not part of the suite)
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
    rng = np.random.default_rng(20240 + SEED)
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
    rng = np.random.default_rng(777 + SEED)
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


def test_random_submissions_state_queries_and_the_static_entry(csr_oracle_lib):
    """The other entry points on the same random graphs: several plans of every sort in ONE submission (rca_expand_many:
    main-stream plans one after the other from the bottom of the output arena, one-launch chains at fixed places at its
    top) against the same plans run one by one; the stand-alone state queries against the oracle; the entry point that
    allocates nothing against rca_expand."""
    import torch
    from k8s_llm_rca_b200 import engine
    rng = np.random.default_rng(4242 + SEED)
    subs = probes = statics = 0
    for g in range(80):
        snap, counts = random_graph(rng)
        if snap.n_rels == 0:
            continue
        orc = csr_oracle_lib.CsrOracle(snap)
        eng = engine.Engine(snap)
        eng.set_graphs(bool(rng.random() < 0.5))
        kinds = [k for k in range(4) if counts[k]]
        # ---- a random submission
        plans, seeds_l, ts_l = [], [], []
        for _ in range(int(rng.integers(2, 12))):
            k0 = int(rng.choice(kinds))
            seeds = snap.nodes_of_kind(k0).astype(np.int32)
            ts = (T0 + rng.integers(-2, 30, size=len(seeds)) * H).astype(np.int64)
            states = [False, True, "mask"][int(rng.integers(0, 3))]
            if rng.random() < 0.5:
                cur, ets, dirs = k0, [], []
                for _step in range(int(rng.integers(1, 5))):
                    opts = [(t, "out", e.dst_kind) for t, e in enumerate(snap.etypes) if e.src_kind == cur] + \
                           [(t, "in", e.src_kind) for t, e in enumerate(snap.etypes) if e.dst_kind == cur]
                    if not opts:
                        break
                    t, d, cur = opts[int(rng.integers(0, len(opts)))]
                    ets.append(t); dirs.append(d)
                if not ets:
                    continue
                plans.append(engine.chain_plan(ets, dirs, with_states=states, count_only=bool(rng.random() < 0.2)))
            else:
                plans.append(engine.varlen_plan(["out", "in", "und"][int(rng.integers(0, 3))], int(rng.integers(1, 5)),
                                                dest_kind=int(rng.integers(-1, 4)), excl_kinds=[4, 5], entity_only=True,
                                                with_states=states, count_only=bool(rng.random() < 0.2),
                                                shortest_only=bool(rng.random() < 0.2)))
            seeds_l.append(seeds); ts_l.append(ts)
        if len(plans) >= 2:
            eng.set_fused_threshold(0 if rng.random() < 0.5 else 1 << 40)
            for rep in range(3):                               # repeated: the second and third run may be graph replays
                many = eng.expand_many(plans, seeds_l, ts_l)
                got = [(list(r.emitted), None if r.count_only else {s: r.rows(s).copy() for s in range(1, r.levels + 1)}) for r in many]
                for (emitted, rows), plan, sd, t in zip(got, plans, seeds_l, ts_l):
                    one = eng.expand(plan, sd, t)
                    assert list(one.emitted) == emitted, (g, rep)
                    if rows is not None:
                        for s in rows:
                            assert np.array_equal(rows[s], one.rows(s)), (g, rep, s)     # same rows in the same order
            subs += 1
        # ---- stand-alone state queries: strict and loose, with and without the label of the state node
        ent = np.concatenate([snap.nodes_of_kind(k) for k in kinds]).astype(np.int32)
        q = ent[rng.integers(0, len(ent), size=64)]
        lo = (T0 + rng.integers(-3, 32, size=64) * H).astype(np.int64)
        hi = lo + rng.integers(0, 4, size=64) * H
        f, c, ids = eng.state_lookup(q, lo)
        wf, wc, wids = orc.states(q, lo)
        assert np.array_equal(f, wf) and np.array_equal(c, wc) and np.array_equal(ids, wids)
        f, c, ids = eng.state_lookup(q, hi, ts_hi=lo)                                    # loose: tmin <= hi and tmax > lo
        wf, wc, wids = orc.states(q, hi, ts_hi=lo)
        assert np.array_equal(f, wf) and np.array_equal(c, wc) and np.array_equal(ids, wids)
        sk = int(rng.integers(6, 10))
        f, c, ids = eng.state_lookup(q, lo, state_kind=sk)
        klo, khi = snap.kind_range(sk)
        _, _, allids = orc.states(q, lo, limit=64)
        for i in range(64):
            keep = [v for v in allids[i].tolist() if klo <= v < khi]
            assert c[i] == len(keep) and ids[i].tolist()[:len(keep[:10])] == keep[:10] and f[i] == (keep[0] if keep else -1)
        probes += 1
        # ---- the entry point that allocates nothing
        k0 = int(rng.choice(kinds))
        seeds = snap.nodes_of_kind(k0).astype(np.int32)
        ts = (T0 + rng.integers(0, 25, size=len(seeds)) * H).astype(np.int64)
        hops = int(rng.integers(1, 6))
        states = [False, True, "mask"][int(rng.integers(0, 3))]
        plan = engine.varlen_plan(["out", "und"][int(rng.integers(0, 2))], hops, dest_kind=int(rng.integers(-1, 4)), excl_kinds=[4, 5],
                                  entity_only=bool(rng.random() < 0.7), with_states=states)
        eng.set_fused_threshold(1 << 40)
        eng.set_stream(torch.cuda.current_stream().cuda_stream)
        d_rows = torch.full((2_000_000,), -9, dtype=torch.int32, device="cuda")
        try:
            st = eng.expand_static(plan, torch.from_numpy(seeds).cuda(), torch.from_numpy(ts).cuda() if states else None, d_rows)
        except engine.RcaError as ex:                          # more partial paths than the fixed scratch holds: refused, never wrong
            assert "RCA_STATIC_MAX_FRONTIER" in str(ex) or "row buffer" in str(ex)
        else:
            ref = eng.expand(plan, seeds, ts if states else None)
            host, off = d_rows.cpu().numpy(), 0
            for s in range(1, hops + 1):
                n, wd = int(st.emitted[s]), int(st.row_len[s])
                assert n == ref.emitted[s]
                if n:
                    assert np.array_equal(host[off:off + n * wd].reshape(n, wd), ref.rows(s))
                off += n * wd
            assert (host[off:off + 32] == -9).all()
            statics += 1
        eng.set_stream(0)
        eng.close()
    assert subs > 50 and probes > 50 and statics > 40

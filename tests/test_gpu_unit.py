"""Unit steps (include/rca_b200.h: rca_set_unit_steps): typed levels with at most one candidate per partial
path run through count_unit / fill_unit.  Rows, their order and every counter must be exactly those of the
general kernels, and the rows must be the oracle's."""
import numpy as np
import pytest

from conftest import rows_to_set

pytestmark = pytest.mark.gpu

ENTITY_KINDS = ["Pod", "ReplicaSet", "Deployment", "Service", "PersistentVolumeClaim", "PersistentVolume",
                "ConfigMap", "Secret", "nfs", "Node"]


@pytest.fixture(scope="module")
def eng(snap_small):
    from k8s_llm_rca_b200 import engine
    e = engine.Engine(snap_small)
    e.set_fused_threshold(0)                 # unit steps belong to the count / scan / fill path
    yield e
    e.close()


def _both(eng, plan, seeds, ts=None):
    """(unit steps on, unit steps off).  With unit steps on the plan also runs without wide entries
    (rca_set_wide_entries): the two ways to hand paths to a final unit level must give the same bytes."""
    out = []
    for on in (True, False):
        eng.set_unit_steps(on)
        res = eng.expand(plan, seeds, ts)
        rows = None if res.count_only else {s: res.rows(s).copy() for s in range(1, res.levels + 1)}
        out.append((res, rows))
        if on:
            eng.set_wide_entries(True)
            wide = eng.expand(plan, seeds, ts)
            assert wide.emitted == res.emitted and wide.frontier == res.frontier and wide.scanned == res.scanned
            if rows is not None:
                for s in rows:
                    assert np.array_equal(wide.rows(s), rows[s]), ("wide entries change level", s)
            eng.set_wide_entries(False)
    eng.set_unit_steps(True)
    return out


def test_every_kind_pair_same_bytes_as_general_kernels(eng, snap_small, csr_oracle_lib):
    from k8s_llm_rca_b200 import engine
    orc = csr_oracle_lib.CsrOracle(snap_small)
    excl = [snap_small.kind_index("Event"), snap_small.kind_index("Namespace")]
    checked = 0
    for direction in ("und", "out", "in"):
        for src in ENTITY_KINDS:
            seeds = snap_small.nodes_of_kind(snap_small.kind_index(src))
            for dst in ENTITY_KINDS:
                for hops in (2, 3, 4, 5, 6):
                    kw = dict(dest_kind=snap_small.kind_index(dst), excl_kinds=excl, entity_only=True)
                    plan = engine.varlen_plan(direction, hops, **kw)
                    (on, rows_on), (off, rows_off) = _both(eng, plan, seeds)
                    tag = (direction, src, dst, hops)
                    assert on.emitted == off.emitted and on.frontier == off.frontier and on.scanned == off.scanned, tag
                    for s in rows_on:
                        assert np.array_equal(rows_on[s], rows_off[s]), tag + (s,)      # same rows in the same order
                    if on.total and hops == 4:
                        ref, _ = orc.varlen(seeds, direction, hops, **kw)
                        assert rows_to_set(rows_on) == rows_to_set(ref), tag
                        checked += 1
                    if hops in (3, 4):
                        # the seed kind declared: the count pass of a unit level may then skip the walk up the path
                        # (it has been shown that nothing before the tail can be of the step's far kind)
                        declared = engine.varlen_plan(direction, hops, pos_kind=[snap_small.kind_index(src)], **kw)
                        (d_on, d_rows_on), (d_off, d_rows_off) = _both(eng, declared, seeds)
                        assert d_on.emitted == d_off.emitted == on.emitted and d_on.frontier == d_off.frontier, tag
                        for s in d_rows_on:
                            assert np.array_equal(d_rows_on[s], d_rows_off[s]), tag + (s, "declared seed kind")
                            assert np.array_equal(np.sort(d_rows_on[s], axis=0), np.sort(rows_on[s], axis=0)), tag + (s,)
    assert checked >= 10


def test_unit_steps_with_states_intermediate_kinds_and_count_only(eng, snap_small, csr_oracle_lib):
    from k8s_llm_rca_b200 import engine
    from k8s_llm_rca_b200.synth import DAY, T_BASE
    orc = csr_oracle_lib.CsrOracle(snap_small)
    k = snap_small.kind_index
    excl = [k("Event"), k("Namespace")]
    seeds = snap_small.nodes_of_kind(k("Pod"))
    ts = (T_BASE + np.random.default_rng(9).integers(0, 30 * DAY, size=len(seeds))).astype(np.int64)
    for kw in (dict(dest_kind=k("Node")), dict(dest_kind=k("Node"), inter_kinds=[k("ReplicaSet"), k("Service")]),
               dict(dest_kind=k("nfs")), dict(dest_kind=k("Deployment"), inter_kinds=[k("ReplicaSet")]),
               dict(dest_kind=k("Node"), pos_kind=[k("Pod"), -1, -1, -1, -1])):
        base = dict(excl_kinds=excl, entity_only=True, **kw)
        plan = engine.varlen_plan("und", 4, with_states=True, **base)
        (on, rows_on), (off, rows_off) = _both(eng, plan, seeds, ts)
        assert on.emitted == off.emitted
        for s in rows_on:                       # full rows: nodes, relationships, state ids, missing mask
            assert np.array_equal(rows_on[s], rows_off[s]), (kw, s)
        ref, _ = orc.varlen(seeds, "und", 4, **base)
        assert rows_to_set({s: rows_on[s][:, :2 * s + 1] for s in rows_on}) == rows_to_set(ref)
        for s, rows in rows_on.items():         # the state columns against the oracle's interval lookup
            if len(rows) == 0:
                continue
            ts_of_seed = dict(zip(seeds.tolist(), ts.tolist()))
            t = np.array([ts_of_seed[v] for v in rows[:, 0].tolist()], dtype=np.int64)
            miss = np.zeros(len(rows), np.int64)
            for d in range(s + 1):
                f, c, _ = orc.states(rows[:, d], t)
                assert np.array_equal(rows[:, 2 * s + 1 + d], f)
                miss |= (c == 0).astype(np.int64) << d
            assert np.array_equal(rows[:, 3 * s + 2].astype(np.int64), miss)
        cnt = engine.varlen_plan("und", 4, count_only=True, **base)
        (c_on, _), (c_off, _) = _both(eng, cnt, seeds)
        assert c_on.emitted == c_off.emitted == on.emitted


def test_unit_steps_survive_arena_growth(snap_small, csr_oracle_lib):
    from k8s_llm_rca_b200 import engine
    orc = csr_oracle_lib.CsrOracle(snap_small)
    k = snap_small.kind_index
    kw = dict(dest_kind=k("Node"), excl_kinds=[k("Event"), k("Namespace")], entity_only=True)
    seeds = snap_small.nodes_of_kind(k("Pod"))
    e = engine.Engine(snap_small)
    e.set_fused_threshold(0)
    e.reserve(64, 256)
    res = e.expand(engine.varlen_plan("und", 4, **kw), seeds)
    got = rows_to_set(res.all_rows())
    assert res.retries >= 1
    ref, _ = orc.varlen(seeds, "und", 4, **kw)
    assert got == rows_to_set(ref)
    e.close()


def test_chain_line_same_bytes_as_chain_walk(snap_10k, csr_oracle_lib):
    """Typed chains whose steps before the last are unit steps run through chain_line; with the unit tables off
    the same plans run through chain_walk.  Same rows, same order, same counters, states included."""
    from k8s_llm_rca_b200 import engine, workloads
    from k8s_llm_rca_b200.synth import DAY, T_BASE
    w = workloads.Config3(snap=snap_10k)
    orc = csr_oracle_lib.CsrOracle(snap_10k)
    e = engine.Engine(snap_10k)
    lo, hi = snap_10k.kind_range(snap_10k.label_index("EVENT"))
    seeds = np.concatenate([np.arange(lo, hi, dtype=np.int32), [-3, snap_10k.n_nodes + 7, 0]]).astype(np.int32)
    ts = (T_BASE + np.random.default_rng(4).integers(0, 30 * DAY, size=len(seeds))).astype(np.int64)
    for name, (ets, dirs, plan) in w.chains.items():
        got = []
        for on in (True, False):
            e.set_unit_steps(on)
            r = e.expand(plan, seeds, ts)
            got.append((r.rows(r.levels).copy(), r.frontier, r.scanned, r.emitted))
        (a, fa, sa, na), (b, fb, sb, nb) = got
        assert np.array_equal(a, b), name
        assert fa == fb and sa == sb and na == nb, name
        L = len(ets)
        ok = seeds[(seeds >= 0) & (seeds < snap_10k.n_nodes)]
        ref, _ = orc.chain(ok, ets, [1 if d == "in" else 0 for d in dirs])
        assert rows_to_set({L: a[:, :2 * L + 1]}) == rows_to_set({L: ref}), name
        cnt = e.expand(engine.chain_plan(ets, dirs, count_only=True), seeds)
        assert cnt.emitted == na
    e.close()


def test_count_only_last_unit_step_taken_by_the_level_before(eng, snap_small):
    """Count-only plans whose last level is one unit step never store the level before it: that level's count
    pass takes the last step itself.  Every counter must equal those of the general kernels (unit steps off)."""
    from k8s_llm_rca_b200 import engine
    excl = [snap_small.kind_index("Event"), snap_small.kind_index("Namespace")]
    nonzero = 0
    for direction in ("und", "out", "in"):
        for src in ENTITY_KINDS:
            seeds = snap_small.nodes_of_kind(snap_small.kind_index(src))
            for dst in ENTITY_KINDS:
                for hops in (2, 3, 4, 5):
                    for inter in ([], [snap_small.kind_index("ReplicaSet"), snap_small.kind_index("PersistentVolume")]):
                        plan = engine.varlen_plan(direction, hops, dest_kind=snap_small.kind_index(dst), excl_kinds=excl,
                                                  inter_kinds=inter, entity_only=True, count_only=True)
                        (on, _), (off, _) = _both(eng, plan, seeds)
                        tag = (direction, src, dst, hops, inter)
                        assert on.emitted == off.emitted and on.scanned == off.scanned and on.frontier == off.frontier, tag
                        nonzero += on.emitted[hops] > 0
    assert nonzero >= 20

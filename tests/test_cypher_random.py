"""Random typed-pattern statements over random small multigraphs: text -> QueryRunner (CPU backend of the tests)
against the brute-force matcher of test_cypher_variants.py.  Directions written or not, labels written or not,
patterns in one clause, comma-separated or in separate clauses, relationship keys as maps or predicates, an extra
predicate on a random variable."""
import numpy as np
import pytest

from oracle_backend import oracle_executor
from test_cypher_variants import brute, expect, keyed

from k8s_llm_rca_b200.props import TableProps
from k8s_llm_rca_b200.snapshot import Kind, Snapshot

KINDS = ["A", "B", "C", "D"]
TYPES = ["T", "U"]
KEYS = ["k1", "k2"]


def random_graph(rng):
    counts = [int(rng.integers(1, 5)) for _ in KINDS]
    start = np.concatenate([[0], np.cumsum(counts)])
    etypes, es, ed, ee = [], [], [], []
    for _ in range(int(rng.integers(3, 9))):
        sk, dk = int(rng.integers(0, 4)), int(rng.integers(0, 4))
        et = (TYPES[int(rng.integers(0, 2))], sk, dk, KEYS[int(rng.integers(0, 2))])
        if et in etypes:
            continue
        etypes.append(et)
        for _ in range(int(rng.integers(1, 7))):
            es.append(int(start[sk] + rng.integers(0, counts[sk])))
            ed.append(int(start[dk] + rng.integers(0, counts[dk])))       # parallel relationships and self-loops happen
            ee.append(len(etypes) - 1)
    kinds = [Kind(k, k, "NativeEntity", False) for k in KINDS]
    snap = Snapshot.build(kinds, counts, etypes, es, ed, ee)
    props = []
    for k, c in zip(KINDS, counts):
        for i in range(c):
            props.append({"kind2": k, "name2": f"{k.lower()}{i}", "rank": str(i)})
    snap.props = TableProps(snap, props)
    return snap


def random_statement(rng, snap):
    """-> (text, clauses for brute(), predicate for brute(), returned columns)."""
    n_hops = int(rng.integers(1, 4))
    # a walk that exists in the schema, so that most statements have rows
    ets = snap.etypes
    hops, cur = [], None
    names = ["n0"]
    for i in range(n_hops):
        cand = [e for e in ets if cur is None or cur in (e.src_kind, e.dst_kind)]
        if not cand:
            break
        e = cand[int(rng.integers(0, len(cand)))]
        if cur is None:
            cur = e.src_kind if rng.random() < 0.5 else e.dst_kind
            first_kind = cur
        forward = (e.src_kind == cur) if e.src_kind != e.dst_kind else bool(rng.random() < 0.5)
        far = e.dst_kind if forward else e.src_kind
        hops.append((e, forward, cur, far))
        cur = far
        names.append(f"n{i + 1}")
    kinds_of = [hops[0][2]] + [h[3] for h in hops]
    labelled = [bool(rng.random() < 0.7) for _ in kinds_of]
    labelled[0] = True                                                   # the seed carries a label
    spec, parts, preds, wheres = [], [], [], []
    for i, (e, forward, near, far) in enumerate(hops):
        style = rng.random()
        direction = "und" if style < 0.3 else ("out" if forward else "in")
        a, b, r = names[i], names[i + 1], f"r{i}"
        la = KINDS[near] if labelled[i] else None
        lb = KINDS[far] if labelled[i + 1] else None
        spec.append((a, la, r, [e.rel_type], direction, b, lb))
        key_in_map = rng.random() < 0.5
        use_key = rng.random() < 0.7
        rel = f"[{r}:{e.rel_type}" + (f" {{key: '{e.key}'}}" if use_key and key_in_map else "") + "]"
        arrow = {"out": f"-{rel}->", "in": f"<-{rel}-", "und": f"-{rel}-"}[direction]
        parts.append((f"({a}{':' + la if la else ''})", arrow, f"({b}{':' + lb if lb else ''})"))
        if use_key:
            preds.append(lambda env, rels, r=r, k=e.key: rels[env[r][1]]["props"]["key"] == k)
            if not key_in_map:
                wheres.append((i, f"{r}.key = '{e.key}'"))
    # an extra predicate on a random variable
    extra = rng.random()
    v = names[int(rng.integers(0, len(names)))]
    if extra < 0.25:
        wheres.append((len(hops) - 1, f"{v}.rank <> '0'"))
        preds.append(lambda env, rels, v=v: env["__nodes"][env[v]]["props"]["rank"] != "0")
    elif extra < 0.5:
        k = KINDS[int(rng.integers(0, 4))]
        wheres.append((len(hops) - 1, f"{v}.kind2 = '{k}'"))
        preds.append(lambda env, rels, v=v, k=k: env["__nodes"][env[v]]["props"]["kind2"] == k)
    elif extra < 0.6:
        wheres.append((len(hops) - 1, f"{names[0]}.name2 <> {names[-1]}.name2"))
        preds.append(lambda env, rels: env["__nodes"][env[names[0]]]["props"]["name2"] != env["__nodes"][env[names[-1]]]["props"]["name2"])
    layout = rng.random()
    if layout < 0.34:                                                    # one long pattern
        text = "MATCH " + parts[0][0] + "".join(arrow + b for (_, arrow, b) in parts)
        clauses = [spec]
        if wheres:
            text += " WHERE " + " AND ".join(w for _, w in wheres)
    elif layout < 0.67:                                                  # comma-separated patterns, one clause
        def again(i):
            return f"({names[i]})" if i else parts[0][0]
        text = "MATCH " + ", ".join((again(i) if i else parts[0][0]) + arrow + b for i, (_, arrow, b) in enumerate(parts))
        clauses = [spec]
        if wheres:
            text += " WHERE " + " AND ".join(w for _, w in wheres)
    else:                                                                # one clause per hop, predicates where they belong
        text, clauses = "", []
        for i, (_, arrow, b) in enumerate(parts):
            text += " MATCH " + (parts[0][0] if i == 0 else f"({names[i]})") + arrow + b
            mine = [w for j, w in wheres if j == i]
            if mine:
                text += " WHERE " + " AND ".join(mine)
            clauses.append([spec[i]])
    cols = names + [f"r{i}" for i in range(len(hops))]
    text += " RETURN " + ", ".join(cols)
    return text.strip(), clauses, preds, cols


@pytest.mark.parametrize("seed", range(6))
def test_random_typed_patterns_against_brute_force(seed):
    rng = np.random.default_rng(1000 + seed)
    with_rows = total = 0
    for _ in range(12):
        snap = random_graph(rng)
        if not snap.etypes:
            continue
        nodes, rels = snap.to_property_graph()
        ex = oracle_executor(snap)
        for _ in range(12):
            text, clauses, preds, cols = random_statement(rng, snap)

            def pred(env):
                env = dict(env)
                env["__nodes"] = nodes
                return all(p(env, rels) for p in preds)
            want = expect(brute(nodes, rels, clauses, pred), cols)
            got = keyed(ex.run_query(text))                              # none of these shapes is declined
            assert got == want, text
            total += 1
            with_rows += bool(want)
    assert total > 80 and with_rows > 30

"""Drop-in for the reference's query executor (common/neo4j_query_executor.py:6-24).

    GpuQueryExecutor(uri, user, password)      # same constructor shape
        .run_query(query, parameters=None)     # -> list[Record], fully materialised
        .close()

`uri` names a snapshot instead of a bolt address (or is an alias of one: register_uri_alias):
    snapshot://<name>            a Snapshot registered in this process with register_snapshot()
    snapshot://<name>#metagraph  the schema graph derived from it (what the reference keeps in its
                                 second Neo4j database, test_all.py:21)
    synth://?objects=N&events=M&seed=S[#metagraph]   a generated graph (synth.generate)
    dump:///abs/path[#metagraph] a node / relationship dump on disk (ingest.load_path: JSON lines, CSV file or CSV directory)

`QueryRunner` is the host logic between Cypher text and the engine: it compiles the text
(cypher.py), resolves labels / kinds / edge types against the snapshot, asks a backend for paths,
chains and state intervals, and shapes the rows into Records.  The product has one backend,
`GpuBackend` (the C ABI); there is no CPU backend in this package.
"""
from __future__ import annotations

import itertools
from urllib.parse import parse_qs, urlparse

import numpy as np

from . import cypher as cy
from . import engine as E
from .graphtypes import Node, Path, Record, Relationship
from .props import ts_to_ms


class ServiceUnavailable(Exception):
    """Raised by the constructor when the uri cannot be resolved - the counterpart of the driver's
    verify_connectivity() failing (neo4j_query_executor.py:9)."""


_REGISTRY: dict = {}
_ALIASES: dict = {}


def register_snapshot(name: str, snap) -> None:
    _REGISTRY[name] = snap


def register_uri_alias(uri: str, target: str) -> None:
    """Let a uri the caller's code has hard-wired - the reference's drivers open
    `bolt://10.1.0.176:7687` (metagraph) and `bolt://10.1.0.174:7687` (state graph), test_all.py:21-22 -
    stand for a snapshot uri: register_uri_alias("bolt://10.1.0.174:7687", "snapshot://cluster")."""
    _ALIASES[uri] = target


def unregister_uri_alias(uri: str) -> None:
    _ALIASES.pop(uri, None)


def resolve_snapshot(uri: str):
    uri = _ALIASES.get(uri, uri)
    u = urlparse(uri)
    if u.scheme == "snapshot":
        name = u.netloc or u.path.lstrip("/")
        if name not in _REGISTRY:
            raise ServiceUnavailable(f"no snapshot registered under {name!r} (uri {uri!r})")
        snap = _REGISTRY[name]
    elif u.scheme == "synth":
        q = parse_qs(u.query)
        key = ("synth", u.query)
        if key not in _REGISTRY:
            from . import synth
            _REGISTRY[key] = synth.generate(int(q.get("objects", ["1000"])[0]),
                                            int(q["events"][0]) if "events" in q else None,
                                            int(q.get("seed", ["20240001"])[0]))
        snap = _REGISTRY[key]
    elif u.scheme == "dump":
        path = (u.netloc + u.path) if u.netloc else u.path
        key = ("dump", path)
        if key not in _REGISTRY:
            from . import ingest
            try:
                _REGISTRY[key] = ingest.load_path(path)
            except (OSError, ValueError, KeyError) as exc:
                raise ServiceUnavailable(f"cannot load the dump at {path!r}: {exc}") from exc
        snap = _REGISTRY[key]
    else:
        raise ServiceUnavailable(f"cannot resolve {uri!r}: expected snapshot://<name>, synth://?objects=.. or dump:///path")
    if u.fragment == "metagraph":
        key = ("meta", id(snap))
        if key not in _REGISTRY:
            _REGISTRY[key] = snap.derive_metagraph()
        snap = _REGISTRY[key]
    elif u.fragment:
        raise ServiceUnavailable(f"unknown fragment #{u.fragment} in {uri!r}")
    return snap


# ------------------------------------------------------------------------------------------ backend
class GpuBackend:
    """Executes plans through the C ABI (engine.Engine).  The only backend of the product."""

    def __init__(self, snap, device: int = 0):
        self.snap = snap
        self.engine = E.Engine(snap, device)
        self._text_loaded = False

    def close(self):
        self.engine.close()

    def varlen(self, seeds, direction, max_hops, min_hops, dest_kind, excl_kinds, inter_kinds, entity_only,
               node_simple, pos_kind):
        plan = E.varlen_plan(direction, max_hops, min_hops=min_hops, dest_kind=dest_kind, excl_kinds=excl_kinds,
                             inter_kinds=inter_kinds, entity_only=entity_only, node_simple=node_simple,
                             pos_kind=pos_kind)
        return self.engine.expand(plan, np.ascontiguousarray(seeds, dtype=np.int32)).all_rows()

    def chain(self, seeds, etypes, dirs):
        res = self.engine.expand(E.chain_plan(etypes, dirs), np.ascontiguousarray(seeds, dtype=np.int32))
        return res.rows(len(etypes))

    def chain_many(self, jobs):
        """[(seeds, etypes, dirs)] -> [rows]: the walks of one statement in one submission (rca_expand_many) - one
        synchronisation however many combinations of edge types the statement's pattern allows."""
        if len(jobs) == 1:
            return [self.chain(*jobs[0])]
        plans = [E.chain_plan(et, d) for _, et, d in jobs]
        res = self.engine.expand_many(plans, [np.ascontiguousarray(sd, dtype=np.int32) for sd, _, _ in jobs])
        return [r.rows(len(et)) for r, (_, et, _) in zip(res, jobs)]

    def states(self, nodes, tmin_le, tmax_gt, state_kind=None):
        return self.engine.state_lookup(nodes, tmin_le, ts_hi=tmax_gt, state_kind=state_kind)

    def contains(self, texts):
        """{index of text: sorted node ids whose `message` contains it} - on the device.
        The statements of one incident ask for the same message more than once (S2, then every S4): the
        answers of the last few hundred texts are kept (the snapshot does not change)."""
        if not self._text_loaded:
            ids, msgs = self.snap.props.messages()
            self.engine.text_load(ids, msgs)
            self._text_loaded = True
            self._contains_memo = {}
        memo = self._contains_memo
        ask = [t for t in dict.fromkeys(texts) if t not in memo]
        if ask:
            hits = self.engine.contains(ask)[1]
            if len(memo) + len(ask) > 1024:
                memo.clear()
            for i, t in enumerate(ask):
                memo[t] = hits.get(i)
        return {i: memo[t] for i, t in enumerate(texts) if memo.get(t) is not None}


# ------------------------------------------------------------------------------------------ runner
class QueryRunner:
    def __init__(self, snap, backend):
        self.snap = snap
        self.backend = backend
        self._label = {k.label: i for i, k in enumerate(snap.kinds)}
        self._kconst = [snap.props.kind_const_props(k) for k in range(snap.n_kinds)]
        self._varlen_cache = {}
        self._chain_rt = {}

    # -- public
    def run(self, query, parameters=None):
        params = dict(parameters or {})
        # A snapshot never changes, so the answer to a metapath statement is a function of its text and parameters:
        # the ladder of find_metapath asks the same few (srcKind, destKind, intermediateKinds) questions for incident
        # after incident.  The rows are kept (entities are immutable views of the snapshot); every call still gets
        # its own list.
        try:
            key = (query, tuple(sorted((k, tuple(v) if isinstance(v, list) else v) for k, v in params.items())))
            hit = self._varlen_cache.get(key)
        except TypeError:
            key, hit = None, None
        if hit is not None:
            return list(hit)
        plan = cy.compile_query(query, params, bind=params)
        if isinstance(plan, cy.NodeScanPlan):
            return self._node_scan(plan, params)
        if isinstance(plan, cy.VarLenPlan):
            hit = self._varlen(plan, params)
            if key is not None:
                if len(self._varlen_cache) >= 256:
                    self._varlen_cache.pop(next(iter(self._varlen_cache)))
                self._varlen_cache[key] = hit
            return list(hit)
        if isinstance(plan, cy.ChainPlan):
            return self._chain(plan, params)
        if isinstance(plan, cy.StatePlan):
            return self._state(plan, params)
        raise cy.CypherSyntaxError("unsupported query")

    # -- helpers
    def _kinds_with(self, prop, value):
        return [k for k in range(self.snap.n_kinds) if self._kconst[k].get(prop) == value]

    def _project(self, ret, envs, params):
        """RETURN items over a list of environments (var -> entity): DISTINCT, ORDER BY, SKIP and LIMIT applied."""
        items, star = ret.items, any(e == ("star",) for (e, _, _) in ret.items)
        keys = [alias or src for (_, alias, src) in items]
        aliases = {alias: i for i, (_, alias, _) in enumerate(items) if alias}
        out, seen, sort_keys = [], set(), []
        stop_at = None if ret.order else (None if ret.limit is None else ret.skip + ret.limit)
        for env in envs:
            if star:                                 # RETURN *: the named variables of the row, in alphabetical order
                names = sorted(v for v in env if not v.startswith(" "))
                items = [(("var", v), None, v) for v in names] + [it for it in ret.items if it[0] != ("star",)]
                keys = [alias or src for (_, alias, src) in items]
                star = False
            vals = tuple(cy.eval_expr(e, env, params) for (e, _, _) in items)
            if ret.distinct:
                try:
                    if vals in seen:
                        continue
                    seen.add(vals)
                except TypeError:
                    pass
            if ret.order:                            # ORDER BY sees the variables of the row and the RETURN aliases
                scope = dict(env)
                scope.update({a: vals[i] for a, i in aliases.items()})
                sort_keys.append(tuple(cy.eval_expr(e, scope, params) for (e, _) in ret.order))
            out.append(Record(keys, vals))
            if stop_at is not None and len(out) >= stop_at:
                break
        if ret.order:
            def rank(v):                             # Cypher's order across types: values, then null last (ascending)
                if v is None:
                    return (9, 0)
                if isinstance(v, bool):
                    return (3, v)
                if isinstance(v, (int, float)):
                    return (2, v)
                if isinstance(v, str):
                    return (1, v)
                return (0, getattr(v, "id", 0))     # nodes / relationships: by id
            idx = list(range(len(out)))
            for j in range(len(ret.order) - 1, -1, -1):          # stable sorts, last key first
                idx.sort(key=lambda i: rank(sort_keys[i][j]), reverse=ret.order[j][1])
            out = [out[i] for i in idx]
        if ret.skip:
            out = out[ret.skip:]
        if ret.limit is not None:
            out = out[:ret.limit]
        return out

    # -- S1 / S6
    def _node_scan(self, plan, params):
        s = self.snap
        if plan.label is not None:
            if plan.label not in self._label:
                return []
            ranges = [s.kind_range(self._label[plan.label])]
            kinds = [self._label[plan.label]]
        else:
            ranges = [(0, s.n_nodes)]
            kinds = list(range(s.n_kinds))
        if plan.id_equals is not None:
            cand = [v for k in kinds for v in s.props.find_by_id(k, plan.id_equals)]
        else:
            cand = itertools.chain.from_iterable(range(lo, hi) for lo, hi in ranges)

        def envs():
            for v in cand:
                env = {plan.var: Node(s, v)}
                if plan.where is None or cy.eval_expr(plan.where, env, params) is True:
                    yield env
        return self._project(plan.ret, envs(), params)

    # -- S3a-d
    def _varlen(self, plan, params):
        s = self.snap
        prop = plan.kind_prop
        src_kinds = self._kinds_with(prop, plan.src_kind)
        dest_kinds = self._kinds_with(prop, plan.dest_kind) if plan.dest_kind is not None else [-1]
        if not src_kinds or not dest_kinds:
            return []
        excl = set()
        if plan.excl_kinds:
            # `not node.kind in [...]` is null - hence not true - for a node without the property
            excl = {k for k in range(s.n_kinds) if self._kconst[k].get(prop) is None or self._kconst[k].get(prop) in plan.excl_kinds}
        inter = [k for k in range(s.n_kinds) if self._kconst[k].get(prop) in plan.inter_kinds] if plan.inter_kinds else []
        if plan.inter_kinds and not inter:
            return []                                  # any(...) over kinds that do not exist is false
        pos_sets = None
        if plan.pos_kinds is not None:
            pos_sets = [None if pk is None else self._kinds_with(prop, pk) for pk in plan.pos_kinds]
            if any(ps is not None and not ps for ps in pos_sets):
                return []
        entity_only = bool(excl) and all(k in excl for k in range(s.n_kinds) if s.kinds[k].is_state)
        seeds = np.concatenate([s.nodes_of_kind(k) for k in src_kinds])
        paths = []
        if pos_sets is not None:                       # fixed-length pattern: one run per combination of kinds
            runs = [(pc[-1], list(pc)) for pc in itertools.product(*[ps if ps is not None else [-1] for ps in pos_sets])]
        elif len(src_kinds) == 1:                      # all seeds are of one kind: say so (it sharpens the pruning)
            runs = [(dk, [src_kinds[0]] + [-1] * plan.max_hops) for dk in dest_kinds]
        else:
            runs = [(dk, None) for dk in dest_kinds]
        for dk, pc in runs:
            levels = self.backend.varlen(seeds, plan.direction, plan.max_hops, plan.min_hops, dk, sorted(excl), inter,
                                         entity_only, plan.node_simple, pc)
            for length, rows in levels.items():
                for r in rows[:, :2 * length + 1].tolist():
                    paths.append((tuple(r[:length + 1]), tuple(r[length + 1:])))
        if len(runs) > 1:
            paths = sorted(set(paths))

        def envs():
            for nodes, rels in paths:
                p = Path([Node(s, v) for v in nodes], [Relationship(s, r) for r in rels])
                env = {}
                if plan.path_var:
                    env[plan.path_var] = p
                if plan.pos_kinds is not None:
                    for name, n in zip(plan.node_vars, p.nodes):
                        if name:
                            env[name] = n
                    for name, r in zip(plan.rel_vars, p.relationships):
                        if name:
                            env[name] = r
                else:
                    if plan.node_vars[0]:
                        env[plan.node_vars[0]] = p.nodes[0]
                    if plan.node_vars[1]:
                        env[plan.node_vars[1]] = p.nodes[-1]
                yield env
        return self._project(plan.ret, envs(), params)

    # -- S2 / S4
    def _seed_nodes(self, plan, kind, params):
        s = self.snap
        lo, hi = s.kind_range(kind)
        if plan.seed_where is None:
            return np.arange(lo, hi, dtype=np.int32)
        # the conjunct the engine answers itself picks the candidates - <seed>.message CONTAINS <text> over the message
        # arena, <seed>.id = <string> over the id table - and whatever else the statement asks of the seed is
        # evaluated on those few, not on every node of the kind
        parts = cy.conjuncts(plan.seed_where)
        cand, rest = None, parts
        for j, w in enumerate(parts):
            if w[0] == "cmp" and w[1] == "contains" and w[2] == ("prop", ("var", plan.seed_var), "message") \
                    and not cy._vars_of(w[3]):
                text = cy.eval_expr(w[3], {}, params)
                if text is None:
                    return np.zeros(0, np.int32)
                cand = sorted(v for v in self.backend.contains([text]).get(0, []) if lo <= v < hi)
            elif w[0] == "cmp" and w[1] == "=" and w[2] == ("prop", ("var", plan.seed_var), "id") and not cy._vars_of(w[3]):
                cand = sorted(s.props.find_by_id(kind, cy.eval_expr(w[3], {}, params)))
            else:
                continue
            rest = parts[:j] + parts[j + 1:]
            break
        if cand is None:
            cand = range(lo, hi)
        if rest:
            cand = [v for v in cand if all(cy.eval_expr(w, {plan.seed_var: Node(s, v)}, params) is True for w in rest)]
        return np.array(list(cand), dtype=np.int32)

    def _chain_shape(self, plan):
        """What of a chain plan depends on the snapshot alone: the kinds each variable may be of, every combination
        (and orientation) of edge types that forms one walk, the row column of each variable.  None: no such walk.
        Kept per plan object - the front end hands out one plan per statement skeleton (cypher._PLANS)."""
        hit = self._chain_rt.get(id(plan))
        if hit is not None and hit[0] is plan:
            return hit[1]
        s = self.snap
        shape = None
        if all(lab is None or lab in self._label for lab in plan.labels.values()):
            # kinds a variable may be of: its label, else what an equality test on a per-kind constant property leaves
            want = {v: ({self._label[lab]} if lab is not None else None) for v, lab in plan.labels.items()}
            for v, tests in plan.hints.items():
                for prop, value in tests:
                    if any(prop in kc for kc in self._kconst):
                        ks = {k for k in range(s.n_kinds) if prop not in self._kconst[k] or self._kconst[k][prop] == value}
                        want[v] = ks if want[v] is None else want[v] & ks
            cands = [[t for t, e in enumerate(s.etypes) if e.rel_type in st.rel_types and (st.key is None or e.key == st.key)]
                     for st in plan.steps]
            combos = []

            def allowed(v, kind):
                return want[v] is None or kind in want[v]

            def rec(i, cur_kind, acc):
                if i == len(plan.steps):
                    combos.append(list(acc))
                    return
                frm, to, d0 = plan.walk[i]
                for t in cands[i]:
                    e = s.etypes[t]
                    for d in (("out", "in") if d0 == "und" else (d0,)):
                        near, far = (e.src_kind, e.dst_kind) if d == "out" else (e.dst_kind, e.src_kind)
                        if (cur_kind is not None and near != cur_kind) or not allowed(frm, near) or not allowed(to, far):
                            continue
                        acc.append((t, near, d))
                        rec(i + 1, far, acc)
                        acc.pop()
            rec(0, None, [])
            col = {plan.seed_var: 0}
            for i, (_, to, _) in enumerate(plan.walk):
                col[to] = i + 1
            shape = (combos, col)
        if len(self._chain_rt) >= 256:
            self._chain_rt.clear()
        self._chain_rt[id(plan)] = (plan, shape)
        return shape

    def _chain(self, plan, params):
        s = self.snap
        shape = self._chain_shape(plan)
        if shape is None:
            return []
        combos, col = shape
        L = len(plan.steps)
        seeds_of = {k: self._seed_nodes(plan, k, params) for k in sorted({c[0][1] for c in combos})}
        if plan.seed_limit is not None and seeds_of:
            # WITH <seed> LIMIT n: which rows survive is unspecified in Cypher; the lowest node ids here
            if len(seeds_of) == 1:
                seeds_of = {k: np.sort(v)[:plan.seed_limit] for k, v in seeds_of.items()}
            else:
                keep = np.sort(np.concatenate(list(seeds_of.values())))[:plan.seed_limit]
                seeds_of = {k: v[np.isin(v, keep)] for k, v in seeds_of.items()}
        all_rows = []
        live = [combo for combo in combos if len(seeds_of[combo[0][1]])]
        jobs = [(seeds_of[combo[0][1]], [t for t, _, _ in combo], [d for _, _, d in combo]) for combo in live]
        if hasattr(self.backend, "chain_many"):
            answers = self.backend.chain_many(jobs) if jobs else []
        else:
            answers = [self.backend.chain(*job) for job in jobs]
        for combo, rows in zip(live, answers):
            for i, (t, _, d) in enumerate(combo):   # an undirected hop matches a self-loop once, not once per orientation
                if plan.walk[i][2] == "und" and d == "in" and s.etypes[t].src_kind == s.etypes[t].dst_kind and len(rows):
                    rows = rows[rows[:, i] != rows[:, i + 1]]
            if len(rows):
                all_rows.append(rows)
        if not all_rows:
            return []
        rows = np.concatenate(all_rows) if len(all_rows) > 1 else all_rows[0]
        for g in plan.unique:                       # no relationship is bound twice within one MATCH clause
            for x in range(len(g)):
                for y in range(x + 1, len(g)):
                    rows = rows[rows[:, L + 1 + g[x]] != rows[:, L + 1 + g[y]]]

        def envs():
            for r in rows.tolist():
                env = {v: Node(s, r[c]) for v, c in col.items()}
                for i, st in enumerate(plan.steps):
                    if st.rel_var:
                        env[st.rel_var] = Relationship(s, r[L + 1 + i])
                if all(cy.eval_expr(c, env, params) is True for c in plan.residual):
                    yield env
        return self._project(plan.ret, envs(), params)

    # -- S5 / S5'
    def _state(self, plan, params):
        s = self.snap
        if plan.entity_label not in self._label:
            return []
        ek = self._label[plan.entity_label]
        if plan.state_label is not None and plan.state_label not in self._label:
            return []
        ents = s.props.find_by_id(ek, plan.entity_id)
        if not ents:
            return []
        try:
            lo, hi = ts_to_ms(plan.lo), ts_to_ms(plan.hi)
        except (ValueError, TypeError) as exc:
            raise cy.CypherSyntaxError(
                f"state query: {exc} - the engine compares timestamps as integers, which equals the reference's "
                f"string comparison only for this fixed-width form") from None
        nodes = np.array(ents, dtype=np.int32)
        # the label of the state node is applied on the device, BEFORE the cut at ten ids per entity: the reference's
        # `RETURN n2 LIMIT 10` (analyze_root_cause.py:76-77) also limits the rows that passed `(n2:KIND)`
        sk = self._label[plan.state_label] if plan.state_label is not None else None
        _, cnt, ids = self.backend.states(nodes, np.full(len(nodes), lo, np.int64), np.full(len(nodes), hi, np.int64), sk)
        found = []                                   # (entity, state) pairs: every state stays with its own entity
        for i in range(len(nodes)):
            for sn in ids[i][:min(int(cnt[i]), ids.shape[1])].tolist():
                found.append((int(nodes[i]), sn))
        found.sort(key=lambda t: (t[1], t[0]))
        envs = ({plan.entity_var: Node(s, en), plan.state_var: Node(s, sn)} for en, sn in found)
        return self._project(plan.ret, envs, params)


# ------------------------------------------------------------------------------------------ executor
class GpuQueryExecutor:
    """Same surface as the reference's Neo4jQueryExecutor; `user` / `password` are accepted and
    ignored (a snapshot has no authentication)."""

    def __init__(self, uri, user=None, password=None, device: int = 0):
        self.uri = uri
        self.snapshot = resolve_snapshot(uri)          # raises ServiceUnavailable: 'verify_connectivity'
        self.backend = GpuBackend(self.snapshot, device)
        self.runner = QueryRunner(self.snapshot, self.backend)

    def close(self):
        if self.backend is not None:
            self.backend.close()
            self.backend = None

    def run_query(self, query, parameters=None):
        if self.backend is None:
            raise ServiceUnavailable("executor is closed")
        return self.runner.run(query, parameters)

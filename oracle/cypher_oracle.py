"""CPU ORACLE - TEST INFRASTRUCTURE, NOT PRODUCT CODE.

A plain-Python restatement, over an explicit property graph, of the Cypher statements the
reference sends to Neo4j on its hot path.  Only tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py may import this package.

PARITY UNPINNED: the arithmetic of this path lives inside the Neo4j server, a third-party
Java program that is not under /root/reference and whose version the reference does not pin
(no requirements file; `from neo4j import GraphDatabase` at common/neo4j_query_executor.py:3).
Neither a Neo4j server, a JVM nor the neo4j Python driver exists in this image, and the
reference holds no golden vectors, fixtures or assertions for this path (its test_*.py are
print-only scripts).  What pins this restatement instead:
  * the query texts themselves (cited per function below) and published openCypher semantics
    - relationship isomorphism inside one MATCH pattern, list predicates all/any/single,
    three-valued logic on missing properties, case-sensitive CONTAINS, len(path) = number of
    relationships;
  * the reference's own Python around the queries, imported unmodified where it runs
    (tests/golden/make_golden.py drives find_metapath / human_generate_cypher_query /
    run_and_filter_query / check_statepath over this oracle and commits the results);
  * an independent enumeration with networkx.all_simple_edge_paths (tests/test_oracle.py);
  * the half-open interval example at check_state/analyze_root_cause.py:62-68.

Graph model: nodes[i] = {'id': i, 'labels': set[str], 'props': dict}, rels[j] = {'id': j,
'type': str, 'start': i, 'end': i, 'props': dict}.  A path is (node-id tuple, rel-id tuple).
"""
from __future__ import annotations

from collections import defaultdict


class PGOracle:
    def __init__(self, nodes, rels):
        self.nodes = nodes
        self.rels = rels
        self.out = defaultdict(list)   # node -> [rel id]
        self.inc = defaultdict(list)
        for r in rels:
            self.out[r["start"]].append(r["id"])
            self.inc[r["end"]].append(r["id"])

    # ------------------------------------------------------------------ helpers
    def prop(self, v, name):
        return self.nodes[v]["props"].get(name)          # absent property == null

    def _hops(self, v, directed):
        """(rel id, other end) for every way one pattern hop can leave v."""
        for r in self.out[v]:
            yield r, self.rels[r]["end"]
        if not directed:
            for r in self.inc[v]:
                if self.rels[r]["start"] != self.rels[r]["end"]:   # a loop is traversed once
                    yield r, self.rels[r]["start"]

    # ------------------------------------------------------------------ S1
    def native_external_kinds(self):
        """find_metapath/find_srckind_metapath_neo4j.py:64-68 - MATCH (n1) WHERE n1.category IN
        ['NativeEntity','ExternalEntity'] RETURN n1.category AS category, n1.kind AS kind."""
        return [(n["props"].get("category"), n["props"].get("kind")) for n in self.nodes
                if n["props"].get("category") in ("NativeEntity", "ExternalEntity")]

    # ------------------------------------------------------------------ S2
    def src_kinds(self, message, limit=5):
        """find_srckind_metapath_neo4j.py:76-84 - (n1:Event)-[s1:HasEvent]->(N1:EVENT) WHERE
        N1.message CONTAINS $message, then (n1:Event)-[r1:ReferInternal]->(n2) WHERE r1.key =
        'involvedObject_uid', RETURN DISTINCT n2.kind2 LIMIT 5.  Row order is unspecified in
        Cypher; here: first appearance scanning relationships by id."""
        seen = []
        for s1 in self.rels:
            if s1["type"] != "HasEvent":
                continue
            n1, N1 = s1["start"], s1["end"]
            if "Event" not in self.nodes[n1]["labels"] or "EVENT" not in self.nodes[N1]["labels"]:
                continue
            msg = self.prop(N1, "message")
            if msg is None or message not in msg:
                continue
            for r1 in self.out[n1]:
                rr = self.rels[r1]
                if rr["type"] == "ReferInternal" and rr["props"].get("key") == "involvedObject_uid":
                    k = self.prop(rr["end"], "kind2")
                    if k not in seen:
                        seen.append(k)
        return seen[:limit]

    # ------------------------------------------------------------------ S3a / S3b, literal form
    def varlen_literal(self, src_kind, dest_kind, inter_kinds, directed, max_hops=3, min_hops=1,
                       kind_prop="kind", excluded=("Event", "Namespace"), rel_ok=None):
        """find_srckind_metapath_neo4j.py:95-116, evaluated the way the text reads: enumerate every
        walk of min..max hops that repeats no relationship (Cypher matches one pattern without
        reusing a relationship), THEN apply the WHERE clause to the finished path.  Exponential -
        tiny graphs only; varlen_paths() below prunes and must agree with this."""
        res = []
        starts = [n["id"] for n in self.nodes if self.prop(n["id"], kind_prop) == src_kind]

        def where(pn):
            if self.prop(pn[-1], kind_prop) != dest_kind:
                return False
            # all(node in nodes(path) WHERE single(x in nodes(path) WHERE x = node))
            if not all(sum(1 for x in pn if x == node) == 1 for node in pn):
                return False
            # all(node in nodes(path) WHERE not node.kind in [...]) ; null kind -> null -> not true
            for node in pn:
                k = self.prop(node, kind_prop)
                if k is None or k in excluded:
                    return False
            if inter_kinds is None or len(inter_kinds) == 0:
                return True
            return any(self.prop(node, kind_prop) in inter_kinds for node in pn[1:-1])

        def walk(pn, pr):
            if len(pr) >= min_hops and where(pn):
                res.append((tuple(pn), tuple(pr)))
            if len(pr) == max_hops:
                return
            for r, w in self._hops(pn[-1], directed):
                if r in pr or (rel_ok is not None and not rel_ok(self.rels[r])):
                    continue
                pn.append(w)
                pr.append(r)
                walk(pn, pr)
                pn.pop()
                pr.pop()

        for s in starts:
            walk([s], [])
        return sorted(res)

    # ------------------------------------------------------------------ S3a / S3b, pruned form
    def varlen_paths(self, src_kind, dest_kind, inter_kinds, directed, max_hops=3, min_hops=1,
                     kind_prop="kind", excluded=("Event", "Namespace"), rel_ok=None, seeds=None):
        """Same result set as varlen_literal.  The two all(...) predicates are prefix-closed (a
        path that violates them cannot be repaired by extending it), so they are tested while
        walking; the destination and any(...) tests stay at emission."""
        res = []
        if seeds is None:
            seeds = [n["id"] for n in self.nodes if self.prop(n["id"], kind_prop) == src_kind]

        def bad(v):
            k = self.prop(v, kind_prop)
            return k is None or k in excluded

        def walk(pn, pr):
            if len(pr) >= min_hops and self.prop(pn[-1], kind_prop) == dest_kind:
                if not inter_kinds or any(self.prop(x, kind_prop) in inter_kinds for x in pn[1:-1]):
                    res.append((tuple(pn), tuple(pr)))
            if len(pr) == max_hops:
                return
            for r, w in self._hops(pn[-1], directed):
                if w in pn or bad(w) or (rel_ok is not None and not rel_ok(self.rels[r])):
                    continue
                pn.append(w)
                pr.append(r)
                walk(pn, pr)
                pn.pop()
                pr.pop()

        for s in seeds:
            if not bad(s):
                walk([s], [])
        return sorted(res)

    # ------------------------------------------------------------------ S3c / S3d
    def single_hop(self, src_kind, dest_kind, kind_prop="kind"):
        """find_srckind_metapath_neo4j.py:119-123 - (n1)-[r1]-(n2), only the two kind tests."""
        res = []
        for n in self.nodes:
            if self.prop(n["id"], kind_prop) != src_kind:
                continue
            for r, w in self._hops(n["id"], False):
                if self.prop(w, kind_prop) == dest_kind:
                    res.append(((n["id"], w), (r,)))
        return sorted(res)

    def via_kind(self, src_kind, dest_kind, mid_kind="Namespace", kind_prop="kind"):
        """find_srckind_metapath_neo4j.py:125-129 - (n1)-[r1]-(n2)-[r2]-(n3) with n2.kind =
        'Namespace'.  r1 and r2 sit in one MATCH pattern, so r1 != r2; nodes may coincide."""
        res = []
        for n in self.nodes:
            if self.prop(n["id"], kind_prop) != src_kind:
                continue
            for r1, m in self._hops(n["id"], False):
                if self.prop(m, kind_prop) != mid_kind:
                    continue
                for r2, w in self._hops(m, False):
                    if r2 != r1 and self.prop(w, kind_prop) == dest_kind:
                        res.append(((n["id"], m, w), (r1, r2)))
        return sorted(res)

    # ------------------------------------------------------------------ S4
    def chain(self, seed_rows, clauses):
        """generate_query/generate_query.py:241-260 - a sequence of MATCH (a:L1)-[r:T]->(b:L2)
        WHERE r.key = K clauses is a natural join on the shared aliases; labels must hold; a
        relationship may be reused by a LATER clause (uniqueness is per MATCH), and one row is
        produced per distinct binding of all relationship variables.

        seed_rows: list of {alias: node id}; clauses: list of (a, L1, r, T, b, L2, key).
        Returns rows as dicts alias -> id (node aliases and relationship aliases share the dict)."""
        by_type = defaultdict(list)
        for r in self.rels:
            by_type[r["type"]].append(r)
        rows = [dict(x) for x in seed_rows]
        for (a, la, rv, t, b, lb, key) in clauses:
            nxt = []
            for row in rows:
                if a in row:
                    cand = [self.rels[r] for r in self.out[row[a]] if self.rels[r]["type"] == t]
                elif b in row:
                    cand = [self.rels[r] for r in self.inc[row[b]] if self.rels[r]["type"] == t]
                else:
                    cand = by_type[t]
                for r in cand:
                    if key is not None and r["props"].get("key") != key:
                        continue
                    if la is not None and la not in self.nodes[r["start"]]["labels"]:
                        continue
                    if lb is not None and lb not in self.nodes[r["end"]]["labels"]:
                        continue
                    if a in row and row[a] != r["start"]:
                        continue
                    if b in row and row[b] != r["end"]:
                        continue
                    if a == b and r["start"] != r["end"]:
                        continue
                    new = dict(row)
                    new[a], new[b], new[rv] = r["start"], r["end"], r["id"]
                    nxt.append(new)
            rows = nxt
        return rows

    def event_seeds(self, message, label="EVENT", limit=1):
        """generate_query.py:223-227 - MATCH (evt:EVENT) WHERE evt.message CONTAINS <msg> WITH evt
        LIMIT 1.  Which node survives LIMIT 1 is unspecified in Cypher; here: lowest node id."""
        hits = [n["id"] for n in self.nodes if label in n["labels"]
                and n["props"].get("message") is not None and message in n["props"]["message"]]
        return hits[:limit] if limit is not None else hits

    # ------------------------------------------------------------------ S5 / S5' / S6
    def strict_states(self, entity_label, entity_id, ts, limit=10):
        """check_state/analyze_root_cause.py:70-79 - (n1:Kind)-[r1:HasState]->(n2:KIND) WHERE n1.id
        = id AND r1.tmin <= ts AND r1.tmax > ts RETURN n2 LIMIT 10.  The comparisons are on
        STRINGS, as in the query text.  Canonical LIMIT: the lowest STATE node ids."""
        return self._states(entity_label, entity_id, lambda lo, hi: lo <= ts and hi > ts, limit)

    def loose_states(self, entity_label, entity_id, tmin, tmax, limit=10):
        """analyze_root_cause.py:51-60 - r1.tmin <= tmax AND r1.tmax > tmin."""
        return self._states(entity_label, entity_id, lambda lo, hi: lo <= tmax and hi > tmin, limit)

    def _states(self, entity_label, entity_id, pred, limit):
        state_label = entity_label.upper()
        hits = []
        for n in self.nodes:
            if entity_label not in n["labels"] or n["props"].get("id") != entity_id:
                continue
            for r in self.out[n["id"]]:
                rr = self.rels[r]
                if rr["type"] != "HasState" or state_label not in self.nodes[rr["end"]]["labels"]:
                    continue
                lo, hi = rr["props"].get("tmin"), rr["props"].get("tmax")
                if lo is None or hi is None:
                    continue
                if pred(lo, hi):
                    hits.append(rr["end"])
        return sorted(hits)[:limit]

    def entity_by_id(self, label, entity_id, limit=1):
        """analyze_root_cause.py:201-206 - match (n1:Kind) where n1.id = id return n1 limit 1."""
        hits = [n["id"] for n in self.nodes if label in n["labels"] and n["props"].get("id") == entity_id]
        return hits[:limit]

"""Incident replay: the graph half of the reference's end-to-end loop (test_all.py:54-140), with the
LLM answers replaced by recorded ones (BASELINE.json: "stubbed with recorded responses").

Per incident (an event message) the reference does, in this order:
  find_srcKind (S2) -> [LLM: destKind, relevant kinds] -> find_metapath (S3 ladder on the metagraph)
  -> per metapath: instance query (S4) + message filter -> per row: state query per node (S5).

Two ways to run it here:
  * `run_one(message)`   - that call order, one statement at a time through the executors' public
                           `run_query` (what a user of the drop-in gets; per-incident latency);
  * `run_batch(messages)`- the same work for many incidents at once: one CONTAINS scan for all messages,
                           one chain expansion per involved-object edge type for the source kinds, the
                           metapath ladder once per distinct (srcKind, destKind) pair, one instance
                           expansion per metapath over all incidents of the group with the state
                           predicate fused into emission, one message-filter launch for all rows.
Both return, per incident, the same canonical result: srcKind, metapaths, the filtered instance rows
and, per row, the entities whose state is missing at the event's time.
"""
from __future__ import annotations

import contextlib
import io
import time

import numpy as np

from . import engine as E
from .executor import QueryRunner
from .graphtypes import Node, Path, Relationship
from .props import ts_to_ms

# recorded answers of the 'root cause locator' assistant: first keyword found in the message wins
RECORDED_DEST = (("nfs export", "nfs", ["PersistentVolumeClaim", "PersistentVolume"]), ("configmap", "ConfigMap", []),
                 ("secret", "Secret", []), ("persistentvolumeclaim", "PersistentVolumeClaim", []),
                 ("nodes are available", "Node", []))


def recorded_dest(message, src_kind):
    for key, dest, inter in RECORDED_DEST:
        if key in message:
            return dest, inter
    return "Node", []


def _name_key(props):
    if props.get("isNative") == "true":
        return "name2", "kind2"
    if props.get("isAtomic") == "true":
        return "val", "tag"
    return {"nfs": "path", "hostPath": "path", "container": "containerName", "image": "imageName"}[props.get("tag")], "tag"


class IncidentReplayer:
    def __init__(self, state_executor, meta_executor, dest_fn=recorded_dest):
        self.sx, self.mx = state_executor, meta_executor
        self.snap = state_executor.snapshot
        self.meta = meta_executor.snapshot
        self.eng = state_executor.backend.engine
        self.dest_fn = dest_fn
        self._mp_cache = {}
        self._needle_cache, self._chain_cache = {}, {}
        s = self.snap
        L = s.label_index
        self._et_has_event = s.find_etypes("HasEvent", L("Event"), L("EVENT"), "metadata_uid")[0]
        self._et_involved = s.find_etypes("ReferInternal", L("Event"), None, "involvedObject_uid")
        self._ev_lo, self._ev_hi = s.kind_range(L("EVENT"))

    # ------------------------------------------------------------------ one by one (reference call order)
    def run_one(self, message):
        from . import dropin
        dropin.install()
        from check_state.analyze_root_cause import find_strict_states
        from find_metapath.find_srckind_metapath_neo4j import find_metapath, find_srcKind
        from generate_query.generate_query import (extend_metapath_construct_string, human_generate_cypher_query,
                                                   run_and_filter_query)
        out = {"rows": [], "missing": []}
        with contextlib.redirect_stdout(io.StringIO()):
            out["srcKind"] = find_srcKind(self.sx, message)
            dest, inter = self.dest_fn(message, out["srcKind"])
            mps = find_metapath(self.mx, out["srcKind"], dest, inter)
            out["metapaths"] = sorted(p.key() for p in mps)
            for mp in sorted(mps, key=lambda p: p.key()):
                q = human_generate_cypher_query(extend_metapath_construct_string(mp), message)
                for rec in run_and_filter_query(self.sx, q):
                    nodes = [x for x in rec if isinstance(x, Node)]
                    ts = next(n["timestamp"] for n in nodes if n["kind"] == "Event")
                    missing = []
                    for n in nodes:
                        if n["kind"] == "Event" or n["kind2"] == "Event":
                            continue
                        kind = n["kind2"] if n["isNative"] == "true" else n["tag"]
                        if not self.sx.run_query(find_strict_states(kind, n["id"], ts)):
                            missing.append(n.id)
                    out["rows"].append(tuple(x.id for x in rec))
                    out["missing"].append(tuple(missing))
        order = sorted(range(len(out["rows"])), key=lambda i: out["rows"][i])
        out["rows"] = [out["rows"][i] for i in order]
        out["missing"] = [out["missing"][i] for i in order]
        return out

    # ------------------------------------------------------------------ batched
    def _metapaths(self, src, dest, inter):
        key = (src, dest, tuple(inter))
        if key not in self._mp_cache:
            from . import dropin
            dropin.install()
            from find_metapath.find_srckind_metapath_neo4j import find_metapath
            with contextlib.redirect_stdout(io.StringIO()):
                try:
                    self._mp_cache[key] = sorted(find_metapath(self.mx, src, dest, inter), key=lambda p: p.key())
                except ValueError:
                    self._mp_cache[key] = []
        return self._mp_cache[key]

    def run_batch(self, messages):
        """The batch as a list of per-incident dictionaries (the form run_one returns)."""
        return self.run_batch_arrays(messages).as_dicts([k.name for k in self.snap.kinds])

    # per kind: the arena of the names the message filter looks for (generate_query.py:104-129), built once
    def _needles(self, k):
        hit = self._needle_cache.get(k)
        if hit is None:
            s = self.snap
            nk, kk = _name_key(s.props.kind_const_props(k))
            names = s.props.column(k, nk)
            kind_text = s.props.kind_const_props(k).get(kk)
            if kind_text is None:
                kinds = s.props.column(k, kk)
                kind_text = kinds[0] if kinds and all(x == kinds[0] for x in kinds) else None
            none = np.fromiter((x is None for x in names), dtype=bool, count=len(names))
            enc = [(x or "").encode("utf-8") for x in names] + [(kind_text or "").encode("utf-8")]
            off = np.zeros(len(enc) + 1, dtype=np.int64)
            np.cumsum(np.fromiter(map(len, enc), dtype=np.int64, count=len(enc)), out=off[1:])
            blob = np.frombuffer(b"".join(enc), dtype=np.uint8).copy() if off[-1] else np.zeros(1, np.uint8)
            hit = self._needle_cache[k] = (blob, off, none, kind_text is None)
        return hit

    def _chain_of(self, src, mp):
        """(edge types, directions) of the instance query of metapath `mp` from source kind `src`: the two event steps,
        then one step per metagraph relationship in the direction the walk crosses it."""
        key = (src, mp.key())
        hit = self._chain_cache.get(key)
        if hit is None:
            s = self.snap
            ets = [self._et_has_event, s.find_etypes("ReferInternal", s.label_index("Event"), s.label_index(src),
                                                      "involvedObject_uid")[0]]
            dirs = ["in", "out"]
            cur = s.label_index(src)
            for rel in mp.relationships:                          # metagraph relationship -> state-graph edge type
                a, b = s.label_index(rel["srcKind"]), s.label_index(rel["destKind"])
                ets.append(s.find_etypes(rel.type, a, b, rel["key"])[0])
                dirs.append("out" if a == cur else "in")
                cur = b if a == cur else a
            hit = self._chain_cache[key] = (ets, dirs, cur)
        return hit

    def run_batch_arrays(self, messages):
        """The same work for many incidents at once, kept in arrays end to end: one CONTAINS scan for all messages, the
        source kinds from one submission of two-step chains, then per (source, destination) group ONE submission of
        its metapaths' instance queries over the group's events with the state predicate fused in (mask rows), and
        the message filter of the group's rows as one launch over a per-kind name arena built once per snapshot.
        No per-row Python.  -> BatchResult."""
        s, eng = self.snap, self.eng
        n = len(messages)
        t_start = time.perf_counter()
        if not self.sx.backend._text_loaded:
            self.sx.backend.contains([])
        evt = eng.contains_first(messages).astype(np.int32)             # lowest EVENT node per message ('LIMIT 1')
        res = BatchResult(n, evt)
        ok = np.nonzero(evt >= 0)[0]
        if len(ok) == 0:
            return res
        uev = np.unique(evt[ok])                                        # everything below is per distinct event
        # S2 for the whole batch: EVENT <-HasEvent- Event -involvedObject-> X, one chain per edge type, one submission
        plans = [E.chain_plan([self._et_has_event, t], ["in", "out"]) for t in self._et_involved]
        src_kind = np.full(len(uev), -1, np.int32)
        for t, r in zip(self._et_involved, eng.expand_many(plans, [uev] * len(plans))):
            rows = r.rows(2)
            if len(rows) == 0:
                continue
            first = np.unique(rows[:, 0], return_index=True)[0]         # one row per event is enough: the kind is the type's
            at = np.searchsorted(uev, first)
            free = src_kind[at] < 0                                     # first edge type wins, like records[0] of S2
            src_kind[at[free]] = s.etypes[t].dst_kind
        # destination kind per distinct event (recorded answer), groups of events by (source, destination, waypoints)
        pos = np.searchsorted(uev, evt[ok])
        k_inc = src_kind[pos]                                           # source kind of every incident with an event
        res.src_kind[ok] = k_inc
        have = k_inc >= 0
        ev_first, first_at = np.unique(evt[ok][have], return_index=True)
        inc_first = ok[have][first_at]                                  # one incident (its message) per distinct event
        names = [k.name for k in s.kinds]
        ev_group, group_ids = {}, {}
        gid_of_uev = np.full(len(uev), -1, np.int32)
        for e_, i, k in zip(ev_first.tolist(), inc_first.tolist(), src_kind[np.searchsorted(uev, ev_first)].tolist()):
            dest, inter = self.dest_fn(messages[i], names[k])
            g = (names[k], dest, tuple(inter))
            gid = group_ids.setdefault(g, len(group_ids))
            ev_group.setdefault(g, []).append(e_)
        for g, evs in ev_group.items():
            gid_of_uev[np.searchsorted(uev, np.asarray(evs, dtype=np.int32))] = group_ids[g]
        res.groups = list(group_ids)
        res.gid = np.full(n, -1, np.int32)
        res.gid[ok] = gid_of_uev[pos]
        for group, evs in ev_group.items():
            src, dest, inter = group
            mps = self._metapaths(src, dest, list(inter))
            res.metapaths[group] = [p.key() for p in mps]
            seeds = np.array(sorted(evs), dtype=np.int32)
            ts_ms = s.props.event_times_ms(seeds)
            chains = [self._chain_of(src, mp) for mp in mps]
            plans = [E.chain_plan(ets, dirs, with_states="mask") for ets, dirs, _ in chains]
            outs = eng.expand_many(plans, [seeds] * len(plans), [ts_ms] * len(plans)) if plans else []
            for mp, (ets, dirs, dk), r in zip(mps, chains, outs):
                L = len(ets)
                rows = r.rows(L)
                if len(rows) == 0:
                    continue
                # message filter (generate_query.py:104-129): the destination's name, or its kind, occurs in the message
                blob, off, none, no_kind = self._needles(dk)
                d_local = rows[:, L] - int(s.kind_start[dk])
                if no_kind or none[d_local].any():                      # the reference's `None in message`
                    raise TypeError("'in <string>' requires string as left operand, not NoneType")
                keep = eng.match_pairs_idx(eng.message_index(rows[:, 0]), d_local, np.full(len(rows), len(off) - 2, np.int32),
                                           blob, off)
                rows = rows[keep]
                if len(rows) == 0:
                    continue
                # column order of the instance query's RETURN: evt, r1, n1, r2, n2, ... (aliases by kind)
                perm = [0] + [c for k in range(1, L + 1) for c in (L + k, k)]
                res.blocks.append(Block(group, mp.key(), L, rows[:, 0].copy(), np.ascontiguousarray(rows[:, perm]),
                                        rows[:, 2 * L + 1].copy()))
            res.group_done_ms[group] = (time.perf_counter() - t_start) * 1e3
        res.finish()
        return res


class Block:
    """Filtered instance rows of one (group, metapath): `evt` the seed EVENT of each row, `rows` in the RETURN order of
    the instance query (evt, r1, n1, r2, n2, ...), `miss` the missing-state mask over path positions."""
    __slots__ = ("group", "mp_key", "L", "evt", "rows", "miss")

    def __init__(self, group, mp_key, L, evt, rows, miss):
        self.group, self.mp_key, self.L, self.evt, self.rows, self.miss = group, mp_key, L, evt, rows, miss


class BatchResult:
    def __init__(self, n, evt):
        self.n, self.evt = n, evt
        self.src_kind = np.full(n, -1, np.int32)      # kind index of the involved object, -1: no such event / object
        self.groups = []                  # (source kind, destination kind, waypoints), by group id
        self.gid = np.full(n, -1, np.int32)
        self.metapaths = {}               # group -> metapath keys
        self.group_done_ms = {}
        self.blocks = []
        self.done_ms = np.zeros(n)
        self.kind_names = None

    def finish(self):
        """Per-incident completion time: when the group its event belongs to was finished."""
        t = np.array([self.group_done_ms.get(g, 0.0) for g in self.groups] + [0.0])
        self.done_ms = t[self.gid]

    @property
    def n_rows(self):
        """Instance rows over all incidents (an event several incidents share counts once per incident)."""
        if not (self.evt >= 0).any():
            return 0
        cnt = np.bincount(self.evt[self.evt >= 0])
        return int(sum(cnt[b.evt].sum() for b in self.blocks))

    def as_dicts(self, kind_names):
        """[{srcKind, metapaths, rows, missing}] per incident - the form IncidentReplayer.run_one returns."""
        by_evt = {}
        for b in self.blocks:
            L = b.L
            for e_, r, m in zip(b.evt.tolist(), b.rows.tolist(), b.miss.tolist()):
                nodes = [r[0]] + r[2::2]                          # path positions 0..L in walk order
                missing = tuple(nodes[k] for k in range(2, L + 1) if (m >> k) & 1)
                by_evt.setdefault(e_, []).append((tuple(r), missing))
        out = []
        for i in range(self.n):
            g = self.groups[self.gid[i]] if self.gid[i] >= 0 else None
            pairs = sorted(by_evt.get(int(self.evt[i]), []), key=lambda t: t[0])
            out.append({"srcKind": kind_names[self.src_kind[i]] if self.src_kind[i] >= 0 else None,
                        "metapaths": list(self.metapaths.get(g, [])) if g else [],
                        "rows": [p[0] for p in pairs], "missing": [p[1] for p in pairs]})
        return out


def latency_profile(replayer, messages, n_single=200):
    """(p50 ms, p99 ms) of run_one over the first n_single messages, and ms per incident of run_batch."""
    lat = []
    for m in messages[:n_single]:
        t0 = time.perf_counter()
        replayer.run_one(m)
        lat.append((time.perf_counter() - t0) * 1e3)
    replayer.run_batch_arrays(messages)          # caches (name arenas, metapaths) are part of a loaded snapshot
    t0 = time.perf_counter()
    replayer.run_batch_arrays(messages)
    batch_ms = (time.perf_counter() - t0) * 1e3
    lat = np.array(lat)
    return {"single_p50_ms": float(np.percentile(lat, 50)), "single_p99_ms": float(np.percentile(lat, 99)),
            "single_n": int(len(lat)), "batch_incidents": len(messages), "batch_total_ms": batch_ms,
            "batch_ms_per_incident": batch_ms / max(len(messages), 1)}

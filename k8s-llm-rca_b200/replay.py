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
        s, eng = self.snap, self.eng
        n = len(messages)
        if not self.sx.backend._text_loaded:
            self.sx.backend.contains([])
        first, _ = eng.contains(list(messages))                     # lowest EVENT node per message ('LIMIT 1')
        evt = first.astype(np.int32)
        results = [{"srcKind": None, "metapaths": [], "rows": [], "missing": []} for _ in range(n)]
        ok = np.nonzero(evt >= 0)[0]
        useeds = np.unique(evt[ok])
        # S2 for the whole batch: EVENT <-HasEvent- Event -involvedObject-> X, one chain per edge type
        obj_of, kind_of = {}, {}
        for t in self._et_involved:
            rows = eng.expand(E.chain_plan([self._et_has_event, t], ["in", "out"]), useeds).rows(2)
            kname = s.kinds[s.etypes[t].dst_kind].name
            for e_, o in zip(rows[:, 0].tolist(), rows[:, 2].tolist()):
                if e_ not in obj_of:                               # first edge type wins, like records[0] of S2
                    obj_of[e_], kind_of[e_] = o, kname
        groups = {}
        for i in ok.tolist():
            e_ = int(evt[i])
            if e_ not in kind_of:
                continue
            src = kind_of[e_]
            dest, inter = self.dest_fn(messages[i], src)
            results[i]["srcKind"] = src
            groups.setdefault((src, dest, tuple(inter)), []).append(i)
        cand = []                                                   # (incident, row, destination name, destination kind)
        for (src, dest, inter), members in groups.items():
            mps = self._metapaths(src, dest, list(inter))
            seeds = np.unique(evt[members])
            ts_ms = np.array([ts_to_ms(s.props.node_props(int(v))["timestamp"]) for v in seeds.tolist()], dtype=np.int64)
            by_evt = {}
            for mp in mps:
                ets = [self._et_has_event, s.find_etypes("ReferInternal", s.label_index("Event"), s.label_index(src),
                                                          "involvedObject_uid")[0]]
                dirs = ["in", "out"]
                cur = s.label_index(src)
                for rel in mp.relationships:                      # metagraph relationship -> state-graph edge type
                    a, b = s.label_index(rel["srcKind"]), s.label_index(rel["destKind"])
                    t = s.find_etypes(rel.type, a, b, rel["key"])[0]
                    ets.append(t)
                    dirs.append("out" if a == cur else "in")
                    cur = b if a == cur else a
                L = len(ets)
                rows = eng.expand(E.chain_plan(ets, dirs, with_states=True), seeds, ts_ms).rows(L)
                # message filter (generate_query.py:104-129): destination named in the message
                dprops = {}
                for r in rows.tolist():
                    e_, d = r[0], r[L]
                    if d not in dprops:
                        p = s.props.node_props(d)
                        nk, kk = _name_key(p)
                        dprops[d] = (p[nk], p[kk])
                    by_evt.setdefault(e_, []).append((mp.key(), r, dprops[d]))
            for i in members:
                results[i]["metapaths"] = [p.key() for p in mps]
                for mpk, r, (dname, dkind) in by_evt.get(int(evt[i]), []):
                    if dname is None or dkind is None:       # the reference's `None in message`
                        raise TypeError("'in <string>' requires string as left operand, not NoneType")
                    cand.append((i, r, dname, dkind))
        # message filter (generate_query.py:104-129) for every row of every incident in one launch: the
        # destination's name or kind must occur in the EVENT's message
        keep = eng.match_pairs([int(evt[c[0]]) for c in cand], [c[2] for c in cand], [c[3] for c in cand])
        for (i, r, _, _), ok in zip(cand, keep.tolist()):
            if not ok:
                continue
            res = results[i]
            L = (len(r) - 2) // 3
            # column order of the instance query's RETURN: evt, r1, n1, r2, n2, ... (aliases by kind)
            cols = [r[0]]
            for k in range(1, L + 1):
                cols += [r[L + k], r[k]]
            miss_mask = r[3 * L + 2]
            missing = tuple(r[k] for k in range(2, L + 1) if (miss_mask >> k) & 1)
            res["rows"].append(tuple(cols))
            res["missing"].append(missing)
        for res in results:
            order = sorted(range(len(res["rows"])), key=lambda k: res["rows"][k])
            res["rows"] = [res["rows"][k] for k in order]
            res["missing"] = [res["missing"][k] for k in order]
        return results


def latency_profile(replayer, messages, n_single=200):
    """(p50 ms, p99 ms) of run_one over the first n_single messages, and ms per incident of run_batch."""
    lat = []
    for m in messages[:n_single]:
        t0 = time.perf_counter()
        replayer.run_one(m)
        lat.append((time.perf_counter() - t0) * 1e3)
    t0 = time.perf_counter()
    replayer.run_batch(messages)
    batch_ms = (time.perf_counter() - t0) * 1e3
    lat = np.array(lat)
    return {"single_p50_ms": float(np.percentile(lat, 50)), "single_p99_ms": float(np.percentile(lat, 99)),
            "single_n": int(len(lat)), "batch_incidents": len(messages), "batch_total_ms": batch_ms,
            "batch_ms_per_incident": batch_ms / max(len(messages), 1)}

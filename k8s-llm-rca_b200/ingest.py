"""Snapshot ingest / export: the on-disk format either side of the path.

Reads a node / relationship dump in the JSON-lines shape of Neo4j's `apoc.export.json.all`
  {"type": "node", "id": "17", "labels": ["Pod"], "properties": {...}}
  {"type": "relationship", "id": "5", "label": "ReferInternal", "start": {"id": "17"}, "end": {"id": "99"},
   "properties": {"key": "spec_nodeName"}}
and builds the kind-major CSR snapshot the engine uploads.  The reference assumes both of its
databases already exist (test_all.py:21-22); its metagraph is derived here from the state graph
(Snapshot.derive_metagraph: one node per entity kind, one relationship per distinct
(relType, srcKind, destKind, key)).

Conventions taken from what the reference's code reads (props.py): a node's kind is its first label;
labels in upper case are STATE kinds, whose `kind` property names the entity kind; entities carry
`kind2` (native) or `tag` (external); HasState relationships carry fixed-width `tmin` / `tmax`.
"""
from __future__ import annotations

import json

import numpy as np

from .props import TableProps, ts_to_ms
from .snapshot import Kind, Snapshot


def _kind_of(labels, props):
    label = labels[0] if labels else "Unlabelled"
    is_state = label.isupper()
    if is_state:
        name = props.get("kind", label.capitalize())
        cat = "NativeState"
    else:
        name = props.get("kind2") or props.get("tag") or label
        cat = props.get("category") or ("ExternalEntity" if props.get("isNative") == "false" else "NativeEntity")
    return label, name, cat, is_state


def load_dump(lines) -> Snapshot:
    """`lines`: a path, or an iterable of JSON strings / dicts."""
    if isinstance(lines, str):
        with open(lines) as f:
            return load_dump(list(f))
    nodes, rels = [], []
    for ln in lines:
        rec = json.loads(ln) if isinstance(ln, (str, bytes)) else ln
        (nodes if rec.get("type") == "node" else rels).append(rec)
    # kinds in order of first appearance; nodes renumbered kind-major, dump order kept inside a kind
    kinds, kidx, members = [], {}, []
    for n in nodes:
        label, name, cat, is_state = _kind_of(n.get("labels", []), n.get("properties", {}))
        if label not in kidx:
            kidx[label] = len(kinds)
            kinds.append(Kind(name, label, cat, is_state))
            members.append([])
        members[kidx[label]].append(n)
    # entity kinds first, then state kinds (keeps STATE ids in one tail range)
    order = sorted(range(len(kinds)), key=lambda k: (kinds[k].is_state, k))
    kinds = [kinds[k] for k in order]
    members = [members[k] for k in order]
    new_id, node_props = {}, []
    for k, ms in enumerate(members):
        for n in ms:
            new_id[str(n["id"])] = len(node_props)
            node_props.append(dict(n.get("properties", {})))
    counts = [len(ms) for ms in members]
    kind_of_new = np.repeat(np.arange(len(kinds)), counts)
    etypes, eidx, es, ed, ee, tmin, tmax, rprops = [], {}, [], [], [], [], [], []
    for r in rels:
        s, d = new_id[str(r["start"]["id"])], new_id[str(r["end"]["id"])]
        p = dict(r.get("properties", {}))
        key = (r.get("label") or r.get("relType"), int(kind_of_new[s]), int(kind_of_new[d]), p.get("key"))
        if key not in eidx:
            eidx[key] = len(etypes)
            etypes.append(key)
        es.append(s)
        ed.append(d)
        ee.append(eidx[key])
        tmin.append(ts_to_ms(p["tmin"]) if "tmin" in p else 0)
        tmax.append(ts_to_ms(p["tmax"]) if "tmax" in p else 0)
        rprops.append({k: v for k, v in p.items() if k not in ("key",)})
    snap = Snapshot.build(kinds, counts, etypes, es, ed, ee, tmin, tmax)
    by_rel = [None] * len(rels)
    for i, rid in enumerate(snap.input_rel_id.tolist()):
        by_rel[rid] = rprops[i]
    snap.props = TableProps(snap, node_props, by_rel)
    return snap


def dump_snapshot(snap, path=None):
    """Inverse of load_dump (nodes in id order, then relationships in id order).  Returns the lines
    when `path` is None."""
    out = []
    for v in range(snap.n_nodes):
        out.append(json.dumps({"type": "node", "id": str(v), "labels": list(snap.props.node_labels(v)),
                               "properties": dict(snap.props.node_props(v))}))
    for r in range(snap.n_rels):
        e = snap.etypes[int(snap.rel_et[r])]
        props = dict(snap.props.rel_props(r))
        props.pop("srcKind", None)
        props.pop("destKind", None)
        out.append(json.dumps({"type": "relationship", "id": str(r), "label": e.rel_type,
                               "start": {"id": str(int(snap.rel_src[r]))}, "end": {"id": str(int(snap.out_dst[r]))},
                               "properties": props}))
    if path is None:
        return out
    with open(path, "w") as f:
        f.write("\n".join(out) + "\n")
    return None

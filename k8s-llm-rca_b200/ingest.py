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

import csv
import glob
import json
import os
import re

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
    new_id, node_props, node_labels = {}, [], []
    for k, ms in enumerate(members):
        for n in ms:
            new_id[str(n["id"])] = len(node_props)
            node_props.append(dict(n.get("properties", {})))
            node_labels.append(tuple(n.get("labels") or (kinds[k].label,)))
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
    # a node's first label is its kind (the range its id lies in); further labels only travel with the node
    extra = any(len(ls) > 1 for ls in node_labels)
    snap.props = TableProps(snap, node_props, by_rel, node_labels=node_labels if extra else None)
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


# ------------------------------------------------------------------------------------------ CSV dumps
# The other shape a Neo4j graph travels in: header-typed CSV files, one or more for nodes and for relationships.
#   neo4j-admin import style:   uid:ID,:LABEL,kind2,isNative:string,...      :START_ID,:END_ID,:TYPE,key,tmin,...
#   apoc.export.csv.all style:  _id,_labels,kind2,...,_start,_end,_type,key  (nodes and relationships in ONE file:
#                               a row with an empty _id is a relationship; labels are ':Pod:Other')
# Every property is kept as the string in the file - the reference compares strings throughout (booleans are the
# strings 'true' / 'false', timestamps fixed-width text: generate_query.py:112-127, analyze_root_cause.py:75) -
# except where the header names a numeric or boolean type (`replicas:int`), which is converted.
_NUMERIC = {"int": int, "long": int, "short": int, "byte": int, "float": float, "double": float}


def _header(cols):
    """-> [(property name or None, role or None, converter or None)] for a header row."""
    out = []
    for c in cols:
        c = c.strip()
        m = re.fullmatch(r"([^:]*):(ID|LABEL|START_ID|END_ID|TYPE|IGNORE)(\([^)]*\))?", c)
        if m:
            out.append((m.group(1) or None, m.group(2), None))
        elif c in ("_id", "_labels", "_start", "_end", "_type"):
            out.append((None, {"_id": "ID", "_labels": "LABEL", "_start": "START_ID", "_end": "END_ID", "_type": "TYPE"}[c], None))
        else:
            name, _, typ = c.partition(":")
            typ = typ.lower()
            conv = _NUMERIC.get(typ) or ((lambda v: v.strip().lower() == "true") if typ == "boolean" else None)
            out.append((name, None, conv))
    return out


def _csv_records(paths):
    """Dump records (the dicts load_dump takes) from header-typed CSV files."""
    if isinstance(paths, str):
        paths = [paths]
    n_rel = 0
    for path in paths:
        with open(path, newline="") as f:
            rd = csv.reader(f)
            try:
                head = _header(next(rd))
            except StopIteration:
                continue
            for row in rd:
                if not row:
                    continue
                role, props = {}, {}
                for (name, what, conv), val in zip(head, row):
                    if what == "IGNORE":
                        continue
                    if what:
                        role[what] = val
                        if what == "ID" and name and val != "":
                            props[name] = val                 # `uid:ID` is a property too, as the import tool keeps it
                    elif val != "":
                        props[name] = conv(val) if conv else val
                if role.get("ID", "") != "" or "START_ID" not in role:
                    labels = [x for x in re.split(r"[;:]", role.get("LABEL", "")) if x]
                    yield {"type": "node", "id": role["ID"], "labels": labels, "properties": props}
                else:
                    yield {"type": "relationship", "id": str(n_rel), "label": role.get("TYPE"),
                           "start": {"id": role["START_ID"]}, "end": {"id": role["END_ID"]}, "properties": props}
                    n_rel += 1


def load_csv(node_files, rel_files=()) -> Snapshot:
    """Header-typed CSV files (either style above) -> snapshot.  Relationship ids are positions in file order."""
    recs = list(_csv_records(node_files)) + list(_csv_records(list(rel_files) if not isinstance(rel_files, str) else rel_files))
    return load_dump(recs)


def dump_csv(snap, directory: str):
    """Write `directory`/nodes.csv and relationships.csv in the neo4j-admin import style (one file each, the
    union of the property names as columns).  Returns the two paths."""
    os.makedirs(directory, exist_ok=True)
    lines = [json.loads(x) for x in dump_snapshot(snap)]
    nodes = [r for r in lines if r["type"] == "node"]
    rels = [r for r in lines if r["type"] == "relationship"]

    def columns(recs):
        seen = {}
        for r in recs:
            for k in r["properties"]:
                seen.setdefault(k, None)
        return list(seen)
    ncols, rcols = columns(nodes), columns(rels)
    npath, rpath = os.path.join(directory, "nodes.csv"), os.path.join(directory, "relationships.csv")
    with open(npath, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow([":ID", ":LABEL"] + ncols)
        for r in nodes:
            w.writerow([r["id"], ";".join(r["labels"])] + [r["properties"].get(c, "") for c in ncols])
    with open(rpath, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow([":START_ID", ":END_ID", ":TYPE"] + rcols)
        for r in rels:
            w.writerow([r["start"]["id"], r["end"]["id"], r["label"]] + [r["properties"].get(c, "") for c in rcols])
    return npath, rpath


def load_path(path: str) -> Snapshot:
    """A dump on disk, by what it is: a JSON-lines file, one CSV file (apoc style, nodes and relationships
    together), or a directory holding `*node*.csv` and `*rel*.csv` files (all of them are read, sorted by name)."""
    if os.path.isdir(path):
        files = sorted(glob.glob(os.path.join(path, "*.csv")))
        nodes = [p for p in files if "node" in os.path.basename(p).lower()]
        rels = [p for p in files if "rel" in os.path.basename(p).lower()]
        other = [p for p in files if p not in nodes and p not in rels]
        if not nodes and not other:
            raise FileNotFoundError(f"no CSV files under {path!r}")
        return load_csv(nodes + other, rels)
    if path.lower().endswith(".csv"):
        return load_csv([path])
    return load_dump(path)

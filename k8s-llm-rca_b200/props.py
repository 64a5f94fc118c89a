"""Property stores: where node / relationship properties live on the host.

The device only ever sees integer ids.  A property store answers, for one node or relationship
id, the property map a Neo4j driver would have hydrated (reference
common/neo4j_query_executor.py:18-21 hydrates every returned node in full; here it is lazy).

Vocabulary follows what the reference reads: `kind`, `category` on metagraph nodes
(find_metapath/find_srckind_metapath_neo4j.py:66-67); `kind2`, `isNative`, `isAtomic`, `tag`,
`id`, `name2`, `val`, `path`, `containerName`, `imageName` on entities
(generate_query/generate_query.py:112-127, check_state/analyze_root_cause.py:97-102,210-219);
`message`, `timestamp` on EVENT (analyze_root_cause.py:86,89); `kind`, `id` on STATE nodes
(:190,232); `key`, `srcKind`, `destKind` on relationships (generate_query.py:55); `tmin`, `tmax`
on HasState (analyze_root_cause.py:75).  Booleans are the strings 'true' / 'false'.
"""
from __future__ import annotations

import numpy as np

_EPOCH = np.datetime64("1970-01-01T00:00:00.000", "ms")
TS_LEN = 23  # 'YYYY-MM-DD HH:MM:SS.mmm'


def ms_to_ts(ms) -> str:
    """int64 milliseconds since the Unix epoch -> 'YYYY-MM-DD HH:MM:SS.mmm' (the format of the
    reference's example lookups, test_check_state.py:41)."""
    s = str(_EPOCH + np.timedelta64(int(ms), "ms"))
    return s.replace("T", " ")


def ts_to_ms(ts: str) -> int:
    """Inverse of ms_to_ts.  The fixed-width check is what makes the reference's *string*
    comparison `r1.tmin <= ts` (analyze_root_cause.py:75) equal to an integer comparison."""
    if len(ts) != TS_LEN or ts[4] != "-" or ts[7] != "-" or ts[10] != " " or ts[13] != ":" \
            or ts[16] != ":" or ts[19] != ".":
        raise ValueError(f"timestamp {ts!r} is not in fixed-width 'YYYY-MM-DD HH:MM:SS.mmm' form")
    return int((np.datetime64(ts.replace(" ", "T"), "ms") - _EPOCH) / np.timedelta64(1, "ms"))


def node_uuid(v: int) -> str:
    return f"{v:08x}-0000-4000-8000-{v:012x}"


def uuid_node(s: str):
    try:
        if len(s) == 36 and s[8:24] == "-0000-4000-8000-":
            v = int(s[:8], 16)
            if int(s[24:], 16) == v:
                return v
    except ValueError:
        pass
    return None


class BaseProps:
    def __init__(self, snap):
        self.snap = snap

    def node_labels(self, v):
        return (self.snap.kinds[int(self.snap.kind_of(v))].label,)

    def rel_props(self, r):
        s = self.snap
        e = s.etypes[int(s.rel_et[r])]
        return {"key": e.key, "srcKind": s.kinds[e.src_kind].name, "destKind": s.kinds[e.dst_kind].name}

    def kind_const_props(self, k):
        """Properties that have the same value on every node of kind `k` (what lets a predicate
        such as `n.kind = 'Pod'` be answered by kind, i.e. by node-id range)."""
        return {}

    def find_by_id(self, kind_idx, id_str):
        """Nodes of kind `kind_idx` whose `id` property equals `id_str`."""
        lo, hi = self.snap.kind_range(kind_idx)
        return [v for v in range(lo, hi) if self.node_props(v).get("id") == id_str]

    def column(self, kind_idx, key):
        """The property `key` of every node of kind `kind_idx`, in node order (None where a node lacks it)."""
        lo, hi = self.snap.kind_range(kind_idx)
        return [self.node_props(v).get(key) for v in range(lo, hi)]

    def event_times_ms(self, nodes):
        """`timestamp` of the given EVENT nodes as int64 milliseconds."""
        return np.array([ts_to_ms(self.node_props(int(v))["timestamp"]) for v in np.asarray(nodes).tolist()], dtype=np.int64)

    def messages(self):
        """(node ids, list of message strings) over every node that has a `message`."""
        ids, msgs = [], []
        for k, kind in enumerate(self.snap.kinds):
            if kind.label != "EVENT":
                continue
            lo, hi = self.snap.kind_range(k)
            for v in range(lo, hi):
                ids.append(v)
                msgs.append(self.node_props(v).get("message"))
        return np.array(ids, dtype=np.int32), msgs


class MetaProps(BaseProps):
    """Metagraph: every node is a kind and carries `kind` and `category`."""

    def node_props(self, v):
        k = self.snap.kinds[int(v)]
        return {"kind": k.name, "category": k.category}

    def kind_const_props(self, k):
        kd = self.snap.kinds[k]
        return {"kind": kd.name, "category": kd.category}


class TableProps(BaseProps):
    """Explicit per-node / per-relationship dictionaries (hand-made graphs in tests, dumps)."""

    def __init__(self, snap, node_props, rel_props=None, node_labels=None, kind_props=None):
        super().__init__(snap)
        self._n = node_props
        self._r = rel_props
        self._l = node_labels
        self._k = kind_props

    def kind_const_props(self, k):
        if self._k is not None:
            return self._k[k]
        lo, hi = self.snap.kind_range(k)
        if lo == hi:
            return {}
        common = dict(self._n[lo])
        for v in range(lo + 1, hi):
            common = {a: b for a, b in common.items() if self._n[v].get(a) == b}
        return common

    def node_props(self, v):
        return self._n[int(v)]

    def node_labels(self, v):
        if self._l is not None:
            return self._l[int(v)]
        return super().node_labels(v)

    def rel_props(self, r):
        base = super().rel_props(r)
        if self._r is not None:
            base.update(self._r[int(r)])
        return base

"""Kind-major CSR / typed-edge snapshot of a property graph.

This is the data format on either side of the hot path: the host builds it once (from the
synthetic generator or from a node/relationship dump), the CUDA engine uploads the integer
arrays to HBM, and the CPU oracles read the same arrays (C oracle) or the property-graph view
derived from them (Python oracle).  Nothing in here traverses the graph.

Numbering rules (they are what makes results comparable across implementations):

* node id   = position in kind-major order: nodes of kind ``k`` occupy
  ``kind_start[k] <= v < kind_start[k+1]``, so ``node.kind == K`` is a range test.
* edge type = dictionary code of ``(relType, srcKind, destKind, key)`` - the tuple the
  reference prints for every metapath step (reference generate_query/generate_query.py:55).
  Types whose endpoints are both entity kinds come first ("class 0"); types that touch a
  STATE kind (HasState, HasEvent) come last ("class 1").
* relationship id = slot in the out-CSR, rows sorted by (src, etype, dst, input order).
* in the (neighbour, rel) pairs of the in- and undirected tables, bit 30 of the rel word is set
  when the hop runs against the relationship's stored direction (REL_REVERSED); ids stay < 2^30.

Every adjacency row therefore lists its entity-entity part first; ``*_rng[v] = (begin,
entity_end)`` and the full row ends at ``*_rng[v+1][0]``.
"""
from __future__ import annotations

import json
import os
from dataclasses import dataclass, field

import numpy as np

I32 = np.int32
I64 = np.int64
REL_REVERSED = 1 << 30
REL_MASK = REL_REVERSED - 1


@dataclass
class Kind:
    name: str            # value of the `kind` property in the metagraph / `kind2` or `tag` on entities
    label: str           # node label ("Pod", "nfs", "POD", "EVENT")
    category: str        # NativeEntity | ExternalEntity | NativeState | ExternalState
    is_state: bool = False


@dataclass
class EdgeType:
    rel_type: str
    src_kind: int
    dst_kind: int
    key: str
    cls: int = 0         # 0: entity-entity, 1: touches a STATE kind


class Snapshot:
    """Integer arrays of one graph plus a property store for lazy hydration."""

    ARRAYS = ("kind_start", "et_src", "et_dst", "et_cls",
              "out_rng", "out_dst", "rel_src", "rel_et",
              "in_rng", "in_adj", "und_rng", "und_adj",
              "typed_out", "typed_out_base", "typed_in", "typed_in_base",
              "st_off", "st_tmin", "st_tmax", "st_node", "st_rel")

    def __init__(self):
        self.kinds: list[Kind] = []
        self.etypes: list[EdgeType] = []
        self.props = None
        for a in self.ARRAYS:
            setattr(self, a, None)

    # ------------------------------------------------------------------ sizes
    @property
    def n_nodes(self) -> int:
        return int(self.kind_start[-1])

    @property
    def n_rels(self) -> int:
        return int(self.out_dst.shape[0])

    @property
    def n_kinds(self) -> int:
        return len(self.kinds)

    @property
    def n_etypes(self) -> int:
        return len(self.etypes)

    def kind_index(self, name: str) -> int:
        for i, k in enumerate(self.kinds):
            if k.name == name and not k.is_state:
                return i
        raise KeyError(name)

    def label_index(self, label: str) -> int:
        for i, k in enumerate(self.kinds):
            if k.label == label:
                return i
        raise KeyError(label)

    def kind_of(self, v):
        return np.searchsorted(self.kind_start, v, side="right") - 1

    def kind_range(self, k: int):
        return int(self.kind_start[k]), int(self.kind_start[k + 1])

    def nodes_of_kind(self, k: int) -> np.ndarray:
        lo, hi = self.kind_range(k)
        return np.arange(lo, hi, dtype=I32)

    def find_etypes(self, rel_type=None, src_kind=None, dst_kind=None, key=None) -> list[int]:
        out = []
        for t, e in enumerate(self.etypes):
            if rel_type is not None and e.rel_type != rel_type:
                continue
            if src_kind is not None and e.src_kind != src_kind:
                continue
            if dst_kind is not None and e.dst_kind != dst_kind:
                continue
            if key is not None and e.key != key:
                continue
            out.append(t)
        return out

    def rel_endpoints(self, r):
        return self.rel_src[r], self.out_dst[r]

    def device_bytes(self) -> int:
        tot = 0
        for a in ("kind_start", "et_src", "et_dst", "out_rng", "out_dst", "in_rng", "in_adj", "und_rng",
                  "und_adj", "typed_out", "typed_in", "st_off", "st_tmin", "st_tmax", "st_node", "rel_et"):
            tot += getattr(self, a).nbytes
        return tot

    # ------------------------------------------------------------------ build
    @classmethod
    def build(cls, kinds, kind_counts, etypes, edge_src, edge_dst, edge_et,
              edge_tmin=None, edge_tmax=None, props=None) -> "Snapshot":
        """Build from an edge list whose node ids are already kind-major.

        `etypes` entries are (relType, srcKindIdx, dstKindIdx, key) in any order; they are
        re-coded into canonical order and `edge_et` is remapped accordingly.
        `edge_tmin/edge_tmax` (int64 ms) matter only for relType 'HasState'.
        Returns the snapshot; `snapshot.input_rel_id[i]` is the relationship id the i-th
        input edge received.
        """
        s = cls()
        s.kinds = [k if isinstance(k, Kind) else Kind(*k) for k in kinds]
        nk = len(s.kinds)
        kc = np.asarray(kind_counts, dtype=I64)
        assert kc.shape == (nk,)
        ks = np.zeros(nk + 1, dtype=I64)
        np.cumsum(kc, out=ks[1:])
        assert ks[-1] < 2**31 - 1, "node ids are int32"
        s.kind_start = ks.astype(I32)
        V = int(ks[-1])

        ets = []
        for e in etypes:
            e = e if isinstance(e, EdgeType) else EdgeType(*e)
            e.cls = 1 if (s.kinds[e.src_kind].is_state or s.kinds[e.dst_kind].is_state) else 0
            ets.append(e)
        order = sorted(range(len(ets)), key=lambda i: (ets[i].cls, ets[i].src_kind, ets[i].dst_kind,
                                                       ets[i].rel_type, ets[i].key))
        remap = np.empty(max(len(ets), 1), dtype=I32)
        for new, old in enumerate(order):
            remap[old] = new
        s.etypes = [ets[i] for i in order]
        nt = len(s.etypes)
        s.et_src = np.array([e.src_kind for e in s.etypes], dtype=I32)
        s.et_dst = np.array([e.dst_kind for e in s.etypes], dtype=I32)
        s.et_cls = np.array([e.cls for e in s.etypes], dtype=I32)

        src = np.asarray(edge_src, dtype=I64)
        dst = np.asarray(edge_dst, dtype=I64)
        et = remap[np.asarray(edge_et, dtype=I64)].astype(I64) if len(src) else np.zeros(0, I64)
        E = len(src)
        assert E < 2**30, "relationship ids are int32 and the undirected table holds 2E entries"
        if E:
            assert src.min() >= 0 and src.max() < V and dst.min() >= 0 and dst.max() < V
            k_src = np.searchsorted(ks, src, side="right") - 1
            k_dst = np.searchsorted(ks, dst, side="right") - 1
            assert np.array_equal(k_src, s.et_src[et]) and np.array_equal(k_dst, s.et_dst[et]), \
                "edge endpoints do not match the kinds of their edge type"
        tmin = np.zeros(E, I64) if edge_tmin is None else np.asarray(edge_tmin, dtype=I64)
        tmax = np.zeros(E, I64) if edge_tmax is None else np.asarray(edge_tmax, dtype=I64)

        # ---- out-CSR; its slot is the relationship id
        perm = np.lexsort((np.arange(E), dst, et, src))
        src, dst, et, tmin, tmax = src[perm], dst[perm], et[perm], tmin[perm], tmax[perm]
        s.input_rel_id = np.empty(E, dtype=I32)
        s.input_rel_id[perm] = np.arange(E, dtype=I32)
        ecls = s.et_cls[et].astype(I64) if E else np.zeros(0, I64)
        s.out_dst = dst.astype(I32)
        s.rel_src = src.astype(I32)
        s.rel_et = et.astype(I32)
        s.out_rng = _ranges(src, ecls, V)

        # ---- in-CSC: (src, rel) pairs, rows sorted by (dst, etype, src, rel)
        rel = np.arange(E, dtype=I64)
        p_in = np.lexsort((rel, src, et, dst))
        s.in_adj = np.stack([src[p_in], rel[p_in] | REL_REVERSED], axis=1).astype(I32) if E else np.zeros((0, 2), I32)
        s.in_rng = _ranges(dst[p_in], ecls[p_in], V)

        # ---- undirected rows: class-0 out, class-0 in, class-1 out, class-1 in; a self-loop is
        #      listed once (its in-copy is dropped)
        keep_in = src != dst
        u_node = np.concatenate([src, dst[keep_in]])
        u_nbr = np.concatenate([dst, src[keep_in]])
        u_rel = np.concatenate([rel, rel[keep_in] | REL_REVERSED])
        u_cls = np.concatenate([ecls, ecls[keep_in]])
        u_part = np.concatenate([np.zeros(E, I64), np.ones(int(keep_in.sum()), I64)])
        u_et = np.concatenate([et, et[keep_in]])
        p_u = np.lexsort((u_rel, u_nbr, u_et, u_part, u_cls, u_node))
        s.und_adj = np.stack([u_nbr[p_u], u_rel[p_u]], axis=1).astype(I32) if E else np.zeros((0, 2), I32)
        s.und_rng = _ranges(u_node[p_u], u_cls[p_u], V)

        # ---- typed ranges: per edge type, (begin, end) slots for every node of its source kind
        #      (out) / destination kind (in)
        s.typed_out, s.typed_out_base = _typed_ranges(et, src, s.et_src, ks, nt)
        s.typed_in, s.typed_in_base = _typed_ranges(et[p_in], dst[p_in], s.et_dst, ks, nt)

        # ---- HasState interval lists, per entity sorted by STATE node id (so that 'LIMIT 10' has a
        #      canonical meaning: the ten lowest STATE node ids)
        is_hs = np.array([e.rel_type == "HasState" for e in s.etypes], dtype=bool)
        hs = np.nonzero(is_hs[et])[0] if E else np.zeros(0, I64)
        p_s = np.lexsort((hs, dst[hs], src[hs]))
        hs = hs[p_s]
        s.st_tmin = tmin[hs].astype(I64)
        s.st_tmax = tmax[hs].astype(I64)
        s.st_node = dst[hs].astype(I32)
        s.st_rel = hs.astype(I32)
        so = np.zeros(V + 1, dtype=I64)
        np.cumsum(np.bincount(src[hs], minlength=V), out=so[1:])
        s.st_off = so.astype(I32)
        s.props = props
        return s

    # ------------------------------------------------------------------ io
    def save(self, path: str) -> None:
        """Write `<path>.npz` (arrays) + `<path>.json` (schema).  Properties are not saved
        here; the property store decides how it persists itself."""
        np.savez(path + ".npz", **{a: getattr(self, a) for a in self.ARRAYS})
        meta = {"kinds": [k.__dict__ for k in self.kinds], "etypes": [e.__dict__ for e in self.etypes]}
        with open(path + ".json", "w") as f:
            json.dump(meta, f)

    @classmethod
    def load(cls, path: str, props=None) -> "Snapshot":
        if not os.path.exists(path + ".npz"):
            raise FileNotFoundError(path + ".npz")
        s = cls()
        with open(path + ".json") as f:
            meta = json.load(f)
        s.kinds = [Kind(**k) for k in meta["kinds"]]
        s.etypes = [EdgeType(**e) for e in meta["etypes"]]
        z = np.load(path + ".npz")
        for a in cls.ARRAYS:
            setattr(s, a, z[a])
        s.props = props
        return s

    # ------------------------------------------------------------------ views
    def to_property_graph(self):
        """Plain-Python property graph for the CPU oracle: (nodes, rels) where
        nodes[v] = {'id': v, 'labels': {..}, 'props': {..}} and
        rels[r]  = {'id': r, 'type': str, 'start': v, 'end': v, 'props': {..}}.
        Small graphs only - every property is hydrated."""
        nodes = []
        for v in range(self.n_nodes):
            nodes.append({"id": v, "labels": set(self.props.node_labels(v)), "props": dict(self.props.node_props(v))})
        rels = []
        for r in range(self.n_rels):
            e = self.etypes[int(self.rel_et[r])]
            rels.append({"id": r, "type": e.rel_type, "start": int(self.rel_src[r]), "end": int(self.out_dst[r]),
                         "props": dict(self.props.rel_props(r))})
        return nodes, rels

    def derive_metagraph(self) -> "Snapshot":
        """One node per entity kind (`category`, `kind`), one relationship per distinct entity-entity
        (relType, srcKind, destKind, key) - the schema graph the reference assumes already exists
        in its 'metagraph' database (reference test_all.py:21)."""
        ent = [i for i, k in enumerate(self.kinds) if not k.is_state]
        idx = {k: j for j, k in enumerate(ent)}
        kinds = [Kind(self.kinds[i].name, self.kinds[i].name, self.kinds[i].category, False) for i in ent]
        ets, es, ed, ee = [], [], [], []
        for e in self.etypes:
            if e.cls != 0:
                continue
            ets.append((e.rel_type, idx[e.src_kind], idx[e.dst_kind], e.key))
            es.append(idx[e.src_kind])
            ed.append(idx[e.dst_kind])
            ee.append(len(ets) - 1)
        from .props import MetaProps
        m = Snapshot.build(kinds, np.ones(len(kinds), I64), ets, es, ed, ee)
        m.props = MetaProps(m)
        return m


def _ranges(row_of, cls_of, V):
    """(begin, entity_end) per row for entries already sorted by (row, class, ...)."""
    n = len(row_of)
    cnt = np.bincount(row_of, minlength=V) if n else np.zeros(V, I64)
    cnt0 = np.bincount(row_of[cls_of == 0], minlength=V) if n else np.zeros(V, I64)
    rng = np.zeros((V + 1, 2), dtype=I64)
    np.cumsum(cnt, out=rng[1:, 0])
    rng[:V, 1] = rng[:V, 0] + cnt0
    rng[V, 1] = rng[V, 0]
    return rng.astype(I32)


def _typed_ranges(et_sorted, row_sorted, et_kind, ks, nt):
    """For entries sorted by (row, etype, ...): per etype a (begin, end) pair for every node of
    the etype's row kind.  Returns (pairs[sum_nodes, 2], base[nt+1])."""
    base = np.zeros(nt + 1, dtype=I64)
    for t in range(nt):
        k = int(et_kind[t])
        base[t + 1] = base[t] + (ks[k + 1] - ks[k])
    pairs = np.zeros((int(base[-1]), 2), dtype=I32)
    if len(et_sorted):
        slots = np.arange(len(et_sorted), dtype=I64)
        order = np.argsort(et_sorted, kind="stable")
        ets = et_sorted[order]
        cut = np.searchsorted(ets, np.arange(nt + 1))
        for t in range(nt):
            sl = slots[order[cut[t]:cut[t + 1]]]
            if len(sl) == 0:
                continue
            k = int(et_kind[t])
            loc = row_sorted[sl] - ks[k]
            nkn = int(ks[k + 1] - ks[k])
            cnt = np.bincount(loc, minlength=nkn)
            first = np.zeros(nkn, dtype=I64)
            # slots of one (row, etype) are contiguous and ascending: the first is the minimum
            starts = np.concatenate([[0], np.cumsum(cnt)[:-1]])
            nz = cnt > 0
            first[nz] = sl[starts[nz]]
            blk = pairs[int(base[t]):int(base[t + 1])]
            blk[:, 0] = first
            blk[:, 1] = first + cnt
    return pairs, base

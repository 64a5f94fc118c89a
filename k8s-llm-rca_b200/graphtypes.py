"""Result types with the surface of the neo4j driver's, as far as the reference uses it.

What the reference touches: `rec['name']`, `rec[int]`, `len(rec)`,
iteration over the record's VALUES (generate_query/generate_query.py:106-110); `x['prop']` on
nodes and relationships, which must give `None` for an absent property because the reference
probes relationships for `'kind'` (generate_query.py:106-107) and nodes for optional keys
(:112-127); `.keys()` (check_state/analyze_root_cause.py:227); `Relationship.type`
(generate_query.py:55); `Path.nodes`, `.relationships`, `len(path)` = number of relationships
(find_metapath/find_srckind_metapath_neo4j.py:153,163-164); `isinstance(x, neo4j.graph.Node)`
(analyze_root_cause.py:85,97).

Entities hold only (snapshot, id); the property map is hydrated on first access.
"""
from __future__ import annotations


class Entity:
    __slots__ = ("_snap", "_id", "_props")

    def __init__(self, snap, ident: int):
        self._snap = snap
        self._id = int(ident)
        self._props = None

    def _load(self):
        raise NotImplementedError

    @property
    def id(self) -> int:
        return self._id

    @property
    def element_id(self) -> str:
        return str(self._id)

    def _p(self) -> dict:
        if self._props is None:
            self._props = self._load()
        return self._props

    def __getitem__(self, name):
        return self._p().get(name)

    def get(self, name, default=None):
        return self._p().get(name, default)

    def __contains__(self, name):
        return name in self._p()

    def keys(self):
        return self._p().keys()

    def values(self):
        return self._p().values()

    def items(self):
        return self._p().items()

    def __iter__(self):
        return iter(self._p())

    def __len__(self):
        return len(self._p())

    def __eq__(self, other):
        return type(other) is type(self) and other._id == self._id and other._snap is self._snap

    def __hash__(self):
        return hash((type(self).__name__, self._id))


class Node(Entity):
    __slots__ = ()

    def _load(self):
        return dict(self._snap.props.node_props(self._id))

    @property
    def labels(self):
        return frozenset(self._snap.props.node_labels(self._id))

    def __repr__(self):
        return f"<Node id={self._id} labels={set(self.labels)} properties={self._p()}>"


class Relationship(Entity):
    __slots__ = ()

    def _load(self):
        return dict(self._snap.props.rel_props(self._id))

    @property
    def type(self) -> str:
        return self._snap.etypes[int(self._snap.rel_et[self._id])].rel_type

    @property
    def start_node(self) -> Node:
        return Node(self._snap, int(self._snap.rel_src[self._id]))

    @property
    def end_node(self) -> Node:
        return Node(self._snap, int(self._snap.out_dst[self._id]))

    @property
    def nodes(self):
        return (self.start_node, self.end_node)

    def __repr__(self):
        return f"<Relationship id={self._id} type={self.type!r} properties={self._p()}>"


class Path:
    """Nodes in walk order; every relationship keeps its stored start/end (so a hop may run against
    it - the 'reversed' steps of reference test_generate_query.py:38)."""
    __slots__ = ("_nodes", "_rels")

    def __init__(self, nodes, rels):
        assert len(nodes) == len(rels) + 1
        self._nodes = tuple(nodes)
        self._rels = tuple(rels)

    @property
    def nodes(self):
        return self._nodes

    @property
    def relationships(self):
        return self._rels

    @property
    def start_node(self):
        return self._nodes[0]

    @property
    def end_node(self):
        return self._nodes[-1]

    def __len__(self):
        return len(self._rels)

    def __iter__(self):
        return iter(self._rels)

    def __eq__(self, other):
        return isinstance(other, Path) and other._nodes == self._nodes and other._rels == self._rels

    def __hash__(self):
        return hash((self._nodes, self._rels))

    def key(self):
        """(node ids, relationship ids): the canonical form results are compared in."""
        return tuple(n.id for n in self._nodes), tuple(r.id for r in self._rels)

    def __repr__(self):
        return f"<Path start={self._nodes[0].id} end={self._nodes[-1].id} size={len(self)}>"


class Record:
    __slots__ = ("_keys", "_values")

    def __init__(self, keys, values):
        self._keys = tuple(keys)
        self._values = tuple(values)

    def keys(self):
        return list(self._keys)

    def values(self):
        return list(self._values)

    def items(self):
        return list(zip(self._keys, self._values))

    def data(self):
        return dict(zip(self._keys, self._values))

    def get(self, key, default=None):
        try:
            return self._values[self._keys.index(key)]
        except ValueError:
            return default

    def __getitem__(self, key):
        if isinstance(key, int):
            return self._values[key]
        if isinstance(key, slice):
            return self._values[key]
        try:
            return self._values[self._keys.index(key)]
        except ValueError:
            raise KeyError(key) from None

    def __len__(self):
        return len(self._values)

    def __iter__(self):
        return iter(self._values)

    def __eq__(self, other):
        return isinstance(other, Record) and other._keys == self._keys and other._values == self._values

    def __hash__(self):
        return hash((self._keys, self._values))

    def __repr__(self):
        return "<Record " + " ".join(f"{k}={v!r}" for k, v in zip(self._keys, self._values)) + ">"

"""TEST HELPER: the executor's backend interface implemented with the CPU oracle (oracle/csr.py).
Lives under tests/ on purpose - the product package has no CPU backend."""
import numpy as np


class OracleBackend:
    def __init__(self, snap):
        from oracle.csr import CsrOracle
        self.snap = snap
        self.orc = CsrOracle(snap)

    def close(self):
        pass

    def varlen(self, seeds, direction, max_hops, min_hops, dest_kind, excl_kinds, inter_kinds, entity_only,
               node_simple, pos_kind):
        levels, _ = self.orc.varlen(seeds, direction, max_hops, min_hops=min_hops, dest_kind=dest_kind,
                                    excl_kinds=excl_kinds, inter_kinds=inter_kinds, entity_only=entity_only,
                                    node_simple=node_simple, pos_kind=pos_kind)
        return levels

    def chain(self, seeds, etypes, dirs):
        rows, _ = self.orc.chain(seeds, etypes, [{"out": 0, "in": 1}[d] for d in dirs])
        return rows

    def states(self, nodes, tmin_le, tmax_gt, state_kind=None):
        if state_kind is None:
            return self.orc.states(nodes, tmin_le, ts_hi=tmax_gt)
        # label first, LIMIT 10 afterwards (reference analyze_root_cause.py:72-77)
        import numpy as np
        _, cnt, ids = self.orc.states(nodes, tmin_le, limit=4096, ts_hi=tmax_gt)
        lo, hi = self.snap.kind_range(state_kind)
        n = len(nodes)
        first, c2, out = np.full(n, -1, np.int32), np.zeros(n, np.int32), np.full((n, 10), -1, np.int32)
        for i in range(n):
            keep = [v for v in ids[i][:int(cnt[i])].tolist() if lo <= v < hi]
            c2[i] = len(keep)
            if keep:
                first[i] = keep[0]
                out[i, :min(10, len(keep))] = keep[:10]
        return first, c2, out

    def contains(self, texts):
        ids, msgs = self.snap.props.messages()
        out = {}
        for i, t in enumerate(texts):
            hit = sorted(int(v) for v, m in zip(ids.tolist(), msgs) if m is not None and t in m)
            if hit:
                out[i] = hit
        return out


def oracle_executor(snap):
    """An object with run_query / close, like the executors, running on the oracle."""
    from k8s_llm_rca_b200.executor import QueryRunner

    class _Ex:
        def __init__(self):
            self.runner = QueryRunner(snap, OracleBackend(snap))

        def run_query(self, query, parameters=None):
            return self.runner.run(query, parameters)

        def close(self):
            pass
    return _Ex()


def record_key(rec):
    """Canonical, comparable form of a Record (entities -> ids, paths -> id tuples)."""
    from k8s_llm_rca_b200.graphtypes import Node, Path, Relationship
    out = []
    for v in rec:
        if isinstance(v, Path):
            out.append(("path",) + v.key())
        elif isinstance(v, Node):
            out.append(("node", v.id))
        elif isinstance(v, Relationship):
            out.append(("rel", v.id))
        else:
            out.append(("val", v))
    return tuple(out)


def oracle_executor_factory(uri, user=None, password=None):
    """For neo4j.set_executor_factory (the shim): the uri resolves as usual, the executor runs on the oracle."""
    from k8s_llm_rca_b200.executor import resolve_snapshot
    ex = oracle_executor(resolve_snapshot(uri))
    ex.snapshot = resolve_snapshot(uri)
    return ex

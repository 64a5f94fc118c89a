"""The benchmark workloads of BASELINE.json, expressed as plans + seeds over a generated graph.

`Config3` is the configuration the headline metric is quoted on: a 1M-object multi-tenant
cluster, 100k seed Pods (every 5th Pod, each with one incident timestamp), per step
  (a) the undirected <=4-hop Pod -> Node enumeration from every seed Pod (the variable-length
      statement of reference find_metapath/find_srckind_metapath_neo4j.py:107-116 applied to the
      instance graph's entity projection), and
  (b) the per-incident instance queries (reference generate_query/generate_query.py:214-260) for
      the metapaths Pod -> {Node, ConfigMap (two edge types), Secret, nfs via PVC <- PV},
both with the state predicate of check_state/analyze_root_cause.py:70-79 fused into emission.
"""
from __future__ import annotations

import numpy as np

from . import engine as E
from . import synth

N_OBJECTS = 1_000_000
N_EVENTS = 108_696            # 92 % of them involve a Pod -> 100,000 seed Pods = every 5th Pod


class Config3:
    name = "config3: 1M-object cluster, 100k seed Pods, undirected <=4-hop Pod->Node + 5 instance queries, fused state filter"

    def __init__(self, snap=None, n_objects=N_OBJECTS, n_events=N_EVENTS, max_hops=4, with_states=True):
        self.snap = snap if snap is not None else synth.generate(n_objects, n_events, synth.config_seed(2))
        s = self.snap
        self.max_hops = max_hops
        K = s.kind_index
        self.excl = [K("Event"), K("Namespace")]
        # position 0 is fixed to Pod: it only restricts the seeds, and lets the schema-level pruning start
        # from the kind the seeds really have
        self.varlen = E.varlen_plan("und", max_hops, dest_kind=K("Node"), excl_kinds=self.excl, entity_only=True,
                                    pos_kind=[K("Pod")], with_states=with_states)
        L = s.label_index

        def et(rel, a, b, key):
            t = s.find_etypes(rel, L(a), L(b), key)
            assert len(t) == 1, (rel, a, b, key)
            return t[0]

        head = [(et("HasEvent", "Event", "EVENT", "metadata_uid"), "in"),
                (et("ReferInternal", "Event", "Pod", "involvedObject_uid"), "out")]
        tails = {
            "Pod->Node": [(et("ReferInternal", "Pod", "Node", "spec_nodeName"), "out")],
            "Pod->ConfigMap(volume)": [(et("ReferInternal", "Pod", "ConfigMap", "spec_volumes_configMap_name"), "out")],
            "Pod->ConfigMap(envFrom)": [(et("ReferInternal", "Pod", "ConfigMap",
                                            "spec_containers_envFrom_configMapRef_name"), "out")],
            "Pod->Secret": [(et("ReferInternal", "Pod", "Secret", "spec_volumes_secret_secretName"), "out")],
            "Pod->PVC<-PV->nfs": [
                (et("ReferInternal", "Pod", "PersistentVolumeClaim", "spec_volumes_persistentVolumeClaim_claimName"), "out"),
                (et("ReferInternal", "PersistentVolume", "PersistentVolumeClaim", "spec_claimRef_uid"), "in"),
                (et("UseExternal", "PersistentVolume", "nfs", "spec_nfs_path"), "out")],
        }
        self.chains = {}
        for name, tail in tails.items():
            steps = head + tail
            self.chains[name] = ([t for t, _ in steps], [d for _, d in steps],
                                 E.chain_plan([t for t, _ in steps], [d for _, d in steps], with_states=with_states))
        self.with_states = with_states
        syn = s.synth
        self.n_inc = int(syn["n_pod_events"])
        self._ev_state_lo = s.kind_range(s.label_index("EVENT"))[0]
        self._pod_lo, self._pod_hi = s.kind_range(K("Pod"))

    def batch(self, rank: int = 0):
        """Incident batch of one rank: (seed Pods, EVENT nodes, timestamps ms).  Every rank gets the
        same number of incidents (weak scaling); rank r's Pods are the (r mod 5)-th residue class
        of the Pod ids, so ranks 0..4 touch disjoint seed sets."""
        s = self.snap
        pods = s.synth["event_pod"][:self.n_inc].astype(np.int64)
        n_pod = self._pod_hi - self._pod_lo
        pods = (pods - self._pod_lo + (rank % 5) + 5 * (rank // 5) * 997) % n_pod + self._pod_lo
        ev = (self._ev_state_lo + np.arange(self.n_inc)).astype(np.int32)
        ts = np.asarray(s.props.ev_ts[:self.n_inc], dtype=np.int64)
        if rank:
            ts = ts + rank * 3_600_000
        return pods.astype(np.int32), ev, ts

    def run(self, eng: E.Engine, pods, ev, ts, fetch=None, batched=True):
        """One step.  `fetch(result)` is called for each expansion while its rows are still resident
        (e2e runs copy them out there).  Returns (instances, [(name, ExpandResult)...]).
        batched: the six plans go to the engine as ONE submission (rca_expand_many); otherwise six calls."""
        if batched:
            t = ts if self.with_states else None
            names = ["varlen"] + list(self.chains)
            plans = [self.varlen] + [c[2] for c in self.chains.values()]
            res = eng.expand_many(plans, [pods] + [ev] * len(self.chains), [t] * len(plans) if t is not None else None)
            if fetch:
                for r in res:
                    fetch(r)
            return sum(r.total for r in res), list(zip(names, res))
        results = []
        total = 0
        r = eng.expand(self.varlen, pods, ts if self.with_states else None)
        if fetch:
            fetch(r)
        results.append(("varlen", r))
        total += r.total
        for name, (_, _, plan) in self.chains.items():
            r = eng.expand(plan, ev, ts if self.with_states else None)
            if fetch:
                fetch(r)
            results.append((name, r))
            total += r.total
        return total, results

    def oracle_run(self, orc, pods, ev, ts, threads=0):
        """The same step on the CPU oracle (callers: tests and bench.py's cpu_baseline leg only):
        the enumeration, the five instance queries, and the state predicate for every node of
        every row.  Returns (instances, {plan name: {length: rows with state columns}})."""
        s = self.snap
        total = 0
        out = {}

        def with_state_cols(rows, length, seeds):
            if not self.with_states or len(rows) == 0:
                return rows
            order = np.argsort(seeds, kind="stable")
            t = ts[order[np.searchsorted(seeds[order], rows[:, 0])]]
            cols = [rows]
            miss = np.zeros(len(rows), np.int64)
            for d in range(length + 1):
                first, cnt, _ = orc.states(rows[:, d], t, limit=1, threads=threads)
                cols.append(first[:, None])
                miss |= (cnt == 0).astype(np.int64) << d
            cols.append(miss.astype(np.int32)[:, None])
            return np.concatenate(cols, axis=1)

        lv, _ = orc.varlen(pods, "und", self.max_hops, dest_kind=s.kind_index("Node"), excl_kinds=self.excl,
                           entity_only=True, threads=threads)
        out["varlen"] = {length: with_state_cols(rows, length, pods) for length, rows in lv.items()}
        total += sum(len(r) for r in lv.values())
        for name, (ets, dirs, _) in self.chains.items():
            rows, _ = orc.chain(ev, ets, [{"out": 0, "in": 1}[d] for d in dirs], threads=threads)
            out[name] = {len(ets): with_state_cols(rows, len(ets), ev)}
            total += len(rows)
        return total, out

    def oracle_step(self, orc, pods, ev, ts, threads=0, pruned=False, keep_rows=False):
        """The same step on the CPU oracle as ONE set of C calls (callers: tests and bench.py's cpu_baseline /
        --impl reference legs only): enumeration + the five instance queries, the state predicate evaluated in C
        for every node of every row - no array glue in the timed call.  pruned: with the kind-level schema pruning
        the engine's plans also use (oracle/csr_oracle.c: orc_reach); the table is built once per workload.
        Returns (instances, {plan name: {length: rows}} or None)."""
        s = self.snap
        kw = dict(dest_kind=s.kind_index("Node"), excl_kinds=self.excl, entity_only=True)
        reach = None
        if pruned:
            if not hasattr(self, "_reach"):
                self._reach = orc.reach("und", self.max_hops, **kw)
            reach = self._reach
        t = ts if self.with_states else None
        out = {} if keep_rows else None
        lv, _, total = orc.varlen_step(pods, t, "und", self.max_hops, reach=reach, threads=threads, keep_rows=keep_rows, **kw)
        if keep_rows:
            out["varlen"] = lv
        for name, (ets, dirs, _) in self.chains.items():
            rows, _ = orc.chain_step(ev, t, ets, [{"out": 0, "in": 1}[d] for d in dirs], threads=threads, keep_rows=keep_rows)
            if keep_rows:
                out[name] = {len(ets): rows}
                total += len(rows)
            else:
                total += rows
        return total, out


def shared_snapshot(n_objects: int, n_events: int, config_index: int, rank: int = 0, barrier=None, cache_dir="/dev/shm"):
    """A generated graph built ONCE per box and shared by the ranks through `cache_dir`: rank 0 generates and
    saves the arrays (Snapshot.save), the others load them after `barrier()`.  Arrays only - no property store:
    the multi-GPU workloads need ids and adjacency, not strings.  A file left by an earlier run of the same
    configuration is reused."""
    import os
    from .snapshot import Snapshot
    path = os.path.join(cache_dir, f"rca_b200_snap_{n_objects}_{n_events}_{synth.config_seed(config_index)}")
    done = path + ".done"
    snap = None
    if rank == 0 and not os.path.exists(done):
        snap = synth.generate(n_objects, n_events, synth.config_seed(config_index))
        try:
            snap.save(path)
            with open(done, "w") as f:
                f.write("ok\n")
        except OSError:
            pass                                   # no room in cache_dir: the other ranks generate their own
    if barrier is not None:
        barrier()
    if snap is None:
        if os.path.exists(done):
            snap = Snapshot.load(path)
        else:
            snap = synth.generate(n_objects, n_events, synth.config_seed(config_index))
    return snap


class Config4:
    """BASELINE config 4's shape: a 10M-object graph, undirected Pod -> Node enumeration of up to 5 hops from a
    sorted seed list that is cut into contiguous, work-balanced shards (parallel.shard_seeds), one per GPU -
    strong scaling: the whole job is fixed, every added GPU takes a share of it.  Two plans:
      count5  <= 5 hops, count only (1.7 k instances per seed: the rows of a million seeds would be 300 GB)
      rows4   <= 4 hops with rows, the rows then gathered to rank 0 (rca_gather_rows) or left where they are."""

    def __init__(self, snap=None, n_objects=10_000_000, n_seeds=1_000_000, rank=0, barrier=None):
        self.snap = snap if snap is not None else shared_snapshot(n_objects, n_objects // 100, 3, rank, barrier)
        s = self.snap
        K = s.kind_index
        pod, node = K("Pod"), K("Node")
        self.excl = [K("Event"), K("Namespace")]
        self.kw = dict(dest_kind=node, excl_kinds=self.excl, entity_only=True)
        pods = s.nodes_of_kind(pod)
        step = max(1, len(pods) // max(1, n_seeds))
        self.seeds = np.ascontiguousarray(pods[::step][:n_seeds], dtype=np.int32)       # sorted ids
        self.count5 = E.varlen_plan("und", 5, pos_kind=[pod], count_only=True, **self.kw)
        self.rows4 = E.varlen_plan("und", 4, pos_kind=[pod], **self.kw)
        self.name = (f"config4 shape: {n_objects / 1e6:g}M-object cluster, {len(self.seeds)} seed Pods in contiguous "
                     f"work-balanced shards, undirected Pod->Node: <=5 hops counted, <=4 hops with rows")

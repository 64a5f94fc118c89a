"""Seeded synthetic multi-tenant Kubernetes state graph.

The reference ships no data and no schema dump; the shape below is inferred from the property
and label names its code reads (see props.py) and the edge types its examples name
(reference generate_query/generate_query.py:51-52,188-190, test_generate_query.py:24-39).
Keys marked (+) are invented in the same naming style to give the graph realistic hubs.

Everything is a pure function of (n_objects, n_events, seed).
"""
from __future__ import annotations

import numpy as np

from .props import BaseProps, ms_to_ts, ts_to_ms, node_uuid, uuid_node
from .snapshot import Kind, Snapshot

# (kind, fraction of n_objects, category, namespaced)
ENTITY_KINDS = [
    ("Pod", 0.50, "NativeEntity", True),
    ("ReplicaSet", 0.06, "NativeEntity", True),
    ("Deployment", 0.05, "NativeEntity", True),
    ("Service", 0.04, "NativeEntity", True),
    ("PersistentVolumeClaim", 0.08, "NativeEntity", True),
    ("PersistentVolume", 0.08, "NativeEntity", False),
    ("ConfigMap", 0.08, "NativeEntity", True),
    ("Secret", 0.06, "NativeEntity", True),
    ("nfs", 0.02, "ExternalEntity", False),
    ("Node", 0.005, "NativeEntity", False),
    ("Namespace", 0.001, "NativeEntity", False),
    ("ResourceQuota", 0.001, "NativeEntity", True),
    ("container", 0.015, "ExternalEntity", False),
    ("image", 0.008, "ExternalEntity", False),
]

T_BASE = ts_to_ms("2024-03-01 00:00:00.000")
T_INF = ts_to_ms("9999-12-31 23:59:59.999")
DAY = 86_400_000


def config_seed(config_index: int) -> int:
    return 20240000 + config_index


def generate(n_objects: int, n_events: int | None = None, seed: int = 20240001) -> Snapshot:
    rng = np.random.default_rng(seed)
    if n_events is None:
        n_events = max(1, n_objects // 10)

    names = [k[0] for k in ENTITY_KINDS] + ["Event"]
    cats = [k[2] for k in ENTITY_KINDS] + ["NativeEntity"]
    nsd = [k[3] for k in ENTITY_KINDS] + [True]
    cnt = [max(1, int(round(k[1] * n_objects))) for k in ENTITY_KINDS] + [n_events]
    KI = {n: i for i, n in enumerate(names)}
    n_ent_kinds = len(names)
    cnt[KI["PersistentVolume"]] = cnt[KI["PersistentVolumeClaim"]]
    T = cnt[KI["Namespace"]]

    # ---- HasState: 1-3 contiguous half-open intervals per entity, 2 % none; 5 % end ("deleted")
    n_ent_total = sum(cnt[:-1])          # every entity but Events has HasState
    m = rng.choice([0, 1, 2, 3], size=n_ent_total, p=[0.02, 0.58, 0.25, 0.15])
    state_cnt_per_kind = []
    off = 0
    for k in range(n_ent_kinds - 1):
        state_cnt_per_kind.append(int(m[off:off + cnt[k]].sum()))
        off += cnt[k]

    kinds = [Kind(names[k], names[k], cats[k], False) for k in range(n_ent_kinds)]
    for k in range(n_ent_kinds - 1):
        kinds.append(Kind(names[k], names[k].upper(), cats[k].replace("Entity", "State"), True))
    kinds.append(Kind("Event", "EVENT", "NativeState", True))
    counts = cnt + state_cnt_per_kind + [n_events]
    kstart = np.zeros(len(counts) + 1, dtype=np.int64)
    np.cumsum(counts, out=kstart[1:])
    SK = lambda k: n_ent_kinds + k       # state kind index of entity kind k
    K_EVENT_STATE = len(kinds) - 1

    def gid(kname, local):
        return kstart[KI[kname]] + local

    # ---- tenancy: objects of a namespaced kind sorted by tenant so each tenant is a local range
    tenant, tstart = {}, {}
    for k, name in enumerate(names):
        if nsd[k] and name != "Event":
            t = np.sort(rng.integers(0, T, size=cnt[k]))
            tenant[name] = t
            tstart[name] = np.searchsorted(t, np.arange(T + 1))

    def pick_in_tenant(dst_kind, tenants):
        lo = tstart[dst_kind][tenants]
        n = tstart[dst_kind][tenants + 1] - lo
        ok = n > 0
        loc = lo + np.floor(rng.random(len(tenants)) * np.maximum(n, 1)).astype(np.int64)
        return loc, ok

    etypes, e_src, e_dst, e_et = [], [], [], []

    def add(rel, sk, dk, key, src_local, dst_local):
        etypes.append((rel, KI[sk] if isinstance(sk, str) else sk, KI[dk] if isinstance(dk, str) else dk, key))
        t = len(etypes) - 1
        s = kstart[etypes[t][1]] + np.asarray(src_local, dtype=np.int64)
        d = kstart[etypes[t][2]] + np.asarray(dst_local, dtype=np.int64)
        e_src.append(s)
        e_dst.append(d)
        e_et.append(np.full(len(s), t, dtype=np.int64))
        return t

    nP = cnt[KI["Pod"]]
    pods = np.arange(nP)
    pten = tenant["Pod"]

    # Pod -> Node (+): uniform, Nodes become ~100-pod hubs
    pod_node = rng.integers(0, cnt[KI["Node"]], size=nP)
    add("ReferInternal", "Pod", "Node", "spec_nodeName", pods, pod_node)
    # Pod -> ReplicaSet -> Deployment (+)
    loc, ok = pick_in_tenant("ReplicaSet", pten)
    ok &= rng.random(nP) < 0.9
    add("ReferInternal", "Pod", "ReplicaSet", "metadata_ownerReferences_uid", pods[ok], loc[ok])
    nR = cnt[KI["ReplicaSet"]]
    loc, ok = pick_in_tenant("Deployment", tenant["ReplicaSet"])
    ok &= rng.random(nR) < 0.9
    add("ReferInternal", "ReplicaSet", "Deployment", "metadata_ownerReferences_uid", np.arange(nR)[ok], loc[ok])
    # Service -> Pod (+): truncated Zipf 1..50 selected pods per service
    nS = cnt[KI["Service"]]
    w = 1.0 / np.arange(1, 51)
    fan = rng.choice(np.arange(1, 51), size=nS, p=w / w.sum())
    svc = np.repeat(np.arange(nS), fan)
    loc, ok = pick_in_tenant("Pod", tenant["Service"][svc])
    pair = np.unique(np.stack([svc[ok], loc[ok]], axis=1), axis=0)
    add("ReferInternal", "Service", "Pod", "spec_selector", pair[:, 0], pair[:, 1])
    # Pod -> PVC ; PV -> PVC ; PV -> nfs   (keys evidenced by the reference's examples)
    loc, ok = pick_in_tenant("PersistentVolumeClaim", pten)
    ok &= rng.random(nP) < 0.3
    pod_pvc = np.where(ok, loc, -1)
    add("ReferInternal", "Pod", "PersistentVolumeClaim", "spec_volumes_persistentVolumeClaim_claimName",
        pods[ok], loc[ok])
    nV = cnt[KI["PersistentVolume"]]
    add("ReferInternal", "PersistentVolume", "PersistentVolumeClaim", "spec_claimRef_uid",
        np.arange(nV), rng.permutation(nV))
    okv = rng.random(nV) < 0.5
    pv_nfs = np.where(okv, rng.integers(0, cnt[KI["nfs"]], size=nV), -1)
    add("UseExternal", "PersistentVolume", "nfs", "spec_nfs_path", np.arange(nV)[okv], pv_nfs[okv])
    # Pod -> ConfigMap (two parallel edge types), Pod -> Secret
    pod_cm = np.full(nP, -1, dtype=np.int64)
    pod_sec = np.full(nP, -1, dtype=np.int64)
    for rep in range(2):
        loc, ok = pick_in_tenant("ConfigMap", pten)
        ok &= rng.random(nP) < (0.7 if rep == 0 else 0.3)
        if rep == 0:
            pod_cm = np.where(ok, loc, -1)
            s0, d0 = pods[ok], loc[ok]
        else:
            add("ReferInternal", "Pod", "ConfigMap", "spec_volumes_configMap_name",
                np.concatenate([s0, pods[ok]]), np.concatenate([d0, loc[ok]]))
    loc, ok = pick_in_tenant("ConfigMap", pten)
    ok &= rng.random(nP) < 0.3
    add("ReferInternal", "Pod", "ConfigMap", "spec_containers_envFrom_configMapRef_name", pods[ok], loc[ok])
    for rep in range(2):
        loc, ok = pick_in_tenant("Secret", pten)
        ok &= rng.random(nP) < (0.7 if rep == 0 else 0.3)
        if rep == 0:
            pod_sec = np.where(ok, loc, -1)
            s0, d0 = pods[ok], loc[ok]
        else:
            add("ReferInternal", "Pod", "Secret", "spec_volumes_secret_secretName",
                np.concatenate([s0, pods[ok]]), np.concatenate([d0, loc[ok]]))
    # Pod -> container -> image (+)
    nC = cnt[KI["container"]]
    add("UseExternal", "Pod", "container", "spec_containers_name", rng.integers(0, nP, size=nC), np.arange(nC))
    add("UseExternal", "container", "image", "spec_containers_image", np.arange(nC),
        rng.integers(0, cnt[KI["image"]], size=nC))

    # ---- events: one Event + EVENT pair per incident; 92 % involve a Pod, the rest a PVC
    n_pod_ev = max(1, int(round(0.92 * n_events))) if n_events > 1 else 1
    n_pvc_ev = n_events - n_pod_ev
    ev_pod = (np.arange(n_pod_ev) * max(1, nP // n_pod_ev)) % nP
    nC_ = cnt[KI["PersistentVolumeClaim"]]
    ev_pvc = (np.arange(n_pvc_ev) * max(1, nC_ // max(n_pvc_ev, 1))) % nC_
    add("ReferInternal", "Event", "Pod", "involvedObject_uid", np.arange(n_pod_ev), ev_pod)
    if n_pvc_ev:
        add("ReferInternal", "Event", "PersistentVolumeClaim", "involvedObject_uid",
            n_pod_ev + np.arange(n_pvc_ev), ev_pvc)
    ev_tenant = np.concatenate([pten[ev_pod], tenant["PersistentVolumeClaim"][ev_pvc]]) if n_pvc_ev else pten[ev_pod]
    # namespaced kind -> Namespace (+)
    for k, name in enumerate(names):
        if name == "Event":
            add("ReferInternal", "Event", "Namespace", "metadata_namespace", np.arange(n_events), ev_tenant)
        elif nsd[k]:
            add("ReferInternal", name, "Namespace", "metadata_namespace", np.arange(cnt[k]), tenant[name])
    etypes.append(("HasEvent", KI["Event"], K_EVENT_STATE, "metadata_uid"))
    t_he = len(etypes) - 1
    e_src.append(kstart[KI["Event"]] + np.arange(n_events))
    e_dst.append(kstart[K_EVENT_STATE] + np.arange(n_events))
    e_et.append(np.full(n_events, t_he, dtype=np.int64))

    # ---- HasState edges with intervals
    tmins, tmaxs = [np.zeros(len(a), np.int64) for a in e_src], [np.zeros(len(a), np.int64) for a in e_src]
    ent_global = np.arange(n_ent_total)            # entity kinds 0..n_ent_kinds-2 are the first ids
    t0 = T_BASE - rng.integers(DAY, 300 * DAY, size=n_ent_total)
    cuts = np.sort(rng.integers(0, 30 * DAY, size=(n_ent_total, 2)), axis=1) + T_BASE
    ended = rng.random(n_ent_total) < 0.05
    t_end = np.where(ended, T_BASE + rng.integers(0, 30 * DAY, size=n_ent_total), T_INF)
    t_end = np.maximum(t_end, cuts[:, 1] + 1)
    bounds = np.stack([t0, cuts[:, 0], cuts[:, 1]], axis=1)     # start of interval j
    state_owner = np.full(int(kstart[-1]), -1, dtype=np.int32)
    off = 0
    for k in range(n_ent_kinds - 1):
        mk = m[off:off + cnt[k]]
        ents = np.repeat(np.arange(cnt[k]), mk)
        j = np.arange(len(ents)) - np.repeat(np.cumsum(mk) - mk, mk)
        ge = off + ents
        lo = bounds[ge, j]
        last = j == (mk[ents] - 1)
        hi = np.where(last, t_end[ge], bounds[ge, np.minimum(j + 1, 2)])
        etypes.append(("HasState", k, SK(k), "metadata_uid"))
        e_src.append(kstart[k] + ents)
        sl = np.arange(len(ents))
        e_dst.append(kstart[SK(k)] + sl)
        e_et.append(np.full(len(ents), len(etypes) - 1, dtype=np.int64))
        tmins.append(lo)
        tmaxs.append(hi)
        state_owner[kstart[SK(k)] + sl] = kstart[k] + ents
        off += cnt[k]
    state_owner[kstart[K_EVENT_STATE]:] = kstart[KI["Event"]] + np.arange(n_events)

    snap = Snapshot.build(kinds, counts, etypes, np.concatenate(e_src), np.concatenate(e_dst),
                          np.concatenate(e_et), np.concatenate(tmins), np.concatenate(tmaxs))

    # ---- messages: each names exactly one object the involved object really points at
    ev_ts = T_BASE + rng.integers(0, 30 * DAY, size=n_events)
    msgs = []
    nN = cnt[KI["Node"]]
    pref = rng.integers(0, 4, size=n_pod_ev)
    for i in range(n_pod_ev):
        p = int(ev_pod[i])
        opts = [("cm", int(pod_cm[p])), ("sec", int(pod_sec[p])), ("pvc", int(pod_pvc[p])), ("node", int(pod_node[p]))]
        opts = opts[pref[i]:] + opts[:pref[i]]
        kind, j = next(o for o in opts if o[1] >= 0)
        if kind == "cm":
            msgs.append(f'MountVolume.SetUp failed for volume "config-vol-{i}" : configmap "configmap-{j}" not found')
        elif kind == "sec":
            msgs.append(f'MountVolume.SetUp failed for volume "secret-vol-{i}" : secret "secret-{j}" not found')
        elif kind == "pvc":
            msgs.append(f'persistentvolumeclaim "persistentvolumeclaim-{j}" not found (incident {i})')
        else:
            msgs.append(f'0/{nN} nodes are available: 1 node(s) had untolerated taint on "node-{j}" (incident {i})')
    for i in range(n_pvc_ev):
        msgs.append(f'failed to provision volume with StorageClass "nfs-client": claim '
                    f'"persistentvolumeclaim-{int(ev_pvc[i])}" is waiting for an nfs export (incident {n_pod_ev + i})')

    snap.props = SynthProps(snap, names, KI, n_ent_kinds, state_owner, ev_ts, msgs)
    snap.synth = {"n_objects": n_objects, "n_events": n_events, "seed": seed,
                  "event_pod": ev_pod.astype(np.int32) + int(kstart[KI["Pod"]]), "n_pod_events": n_pod_ev}
    return snap


class SynthProps(BaseProps):
    """Formula-backed property store of a generated graph: nothing but the event messages and a
    state->owner table is stored; every other property is a function of the node id."""

    def __init__(self, snap, names, KI, n_ent_kinds, state_owner, ev_ts, msgs):
        super().__init__(snap)
        self.names, self.KI, self.n_ent_kinds = names, KI, n_ent_kinds
        self.state_owner, self.ev_ts, self.msgs = state_owner, ev_ts, msgs
        self._st_order = np.argsort(snap.st_rel, kind="stable")
        self._st_rel_sorted = snap.st_rel[self._st_order]

    def node_props(self, v):
        s = self.snap
        v = int(v)
        k = int(s.kind_of(v))
        kind = s.kinds[k]
        j = v - int(s.kind_start[k])
        if not kind.is_state:
            return self._entity(kind, v, j)
        owner = int(self.state_owner[v])
        if kind.label == "EVENT":
            return {"kind": "Event", "id": node_uuid(owner), "message": self.msgs[j],
                    "timestamp": ms_to_ts(self.ev_ts[j]), "type": "Warning", "reason": "Failed"}
        ok = s.kinds[int(s.kind_of(owner))]
        oj = owner - int(s.kind_start[int(s.kind_of(owner))])
        p = {"kind": kind.name, "id": node_uuid(owner),
             "metadata": '{"name": "%s", "generation": %d}' % (self._name(ok, oj), j % 7)}
        if kind.name in ("nfs",):
            p["path"] = f"/exports/vol-{oj}"
            p["server"] = f"10.0.{oj % 250}.{(oj // 250) % 250}"
        elif kind.name == "Secret" or kind.name == "ConfigMap":
            p["data"] = '{"keys": %d}' % (1 + j % 5)
        else:
            p["spec"] = '{"replicas": %d}' % (1 + j % 3)
            p["status"] = '{"phase": "%s"}' % ("Running", "Pending", "Failed")[j % 3]
        return p

    def kind_const_props(self, k):
        kind = self.snap.kinds[k]
        if kind.is_state:
            return {"kind": kind.name}
        if kind.category == "NativeEntity":
            return {"kind2": kind.name, "isNative": "true", "isAtomic": "false"}
        return {"tag": kind.name, "isNative": "false", "isAtomic": "false"}

    @staticmethod
    def _name(kind, j):
        return f"{kind.name.lower()}-{j}"

    def _entity(self, kind, v, j):
        if kind.category == "NativeEntity":
            return {"id": node_uuid(v), "kind2": kind.name, "isNative": "true", "isAtomic": "false",
                    "name2": self._name(kind, j)}
        p = {"id": node_uuid(v), "tag": kind.name, "isNative": "false", "isAtomic": "false"}
        if kind.name == "nfs":
            p["path"] = f"/exports/vol-{j}"
            p["server"] = f"10.0.{j % 250}.{(j // 250) % 250}"
        elif kind.name == "container":
            p["containerName"] = f"container-{j}"
        elif kind.name == "image":
            p["imageName"] = f"registry.local/app-{j}:v1"
        return p

    def column(self, kind_idx, key):
        kind = self.snap.kinds[kind_idx]
        lo, hi = self.snap.kind_range(kind_idx)
        n = hi - lo
        if not kind.is_state:                         # the formulas of _entity, a whole kind at a time
            if key == "name2" and kind.category == "NativeEntity":
                prefix = kind.name.lower()
                return [f"{prefix}-{j}" for j in range(n)]
            if key == "path" and kind.name == "nfs":
                return [f"/exports/vol-{j}" for j in range(n)]
            if key == "containerName" and kind.name == "container":
                return [f"container-{j}" for j in range(n)]
            if key == "imageName" and kind.name == "image":
                return [f"registry.local/app-{j}:v1" for j in range(n)]
        return super().column(kind_idx, key)

    def event_times_ms(self, nodes):
        lo, _ = self.snap.kind_range(len(self.snap.kinds) - 1)
        return np.asarray(self.ev_ts, dtype=np.int64)[np.asarray(nodes, dtype=np.int64) - lo]

    def rel_props(self, r):
        p = super().rel_props(r)
        s = self.snap
        if s.etypes[int(s.rel_et[r])].rel_type == "HasState":
            i = int(np.searchsorted(self._st_rel_sorted, r))
            i = int(self._st_order[i])
            p["tmin"] = ms_to_ts(s.st_tmin[i])
            p["tmax"] = ms_to_ts(s.st_tmax[i])
        return p

    def find_by_id(self, kind_idx, id_str):
        s = self.snap
        v = uuid_node(id_str)
        if v is None or not (0 <= v < s.n_nodes):
            return []
        lo, hi = s.kind_range(kind_idx)
        if not s.kinds[kind_idx].is_state:
            return [v] if lo <= v < hi else []
        return [int(n) for n in np.nonzero(self.state_owner[lo:hi] == v)[0] + lo]

    def messages(self):
        lo, hi = self.snap.kind_range(len(self.snap.kinds) - 1)
        return np.arange(lo, hi, dtype=np.int32), self.msgs

"""The fused state predicate on entities with MANY HasState intervals.  The kernels answer a node from a packed header
(first interval), a packed second record (second and third interval) and, from the fourth interval on, the full
records; the generator only makes one to three intervals per entity, so this hand-made graph has Pods with zero to
seven, overlapping and disjoint, and timestamps that fall into each of them, between them and outside all of them.
Rows with state columns must equal the CPU oracle's (orc_varlen_step), through both kernel families."""
import numpy as np
import pytest

from conftest import sort_rows

pytestmark = pytest.mark.gpu
T0 = 1_700_000_000_000
H = 3_600_000


def _dump():
    from k8s_llm_rca_b200.props import ms_to_ts
    lines, nid, rid = [], 0, 0

    def node(label, **props):
        nonlocal nid
        lines.append({"type": "node", "id": str(nid), "labels": [label], "properties": props})
        nid += 1
        return nid - 1

    def rel(label, a, b, **props):
        nonlocal rid
        lines.append({"type": "relationship", "id": str(rid), "label": label, "start": {"id": str(a)}, "end": {"id": str(b)}, "properties": props})
        rid += 1
    hosts = [node("Node", kind2="Node", isNative="true", id=f"n{j}", name2=f"node-{j}") for j in range(2)]
    pods = []
    for i in range(8):
        p = node("Pod", kind2="Pod", isNative="true", id=f"p{i}", name2=f"pod-{i}")
        pods.append(p)
        rel("ReferInternal", p, hosts[i % 2], key="spec_nodeName")
    for i, p in enumerate(pods):                      # Pod i has i intervals: [k, k+2) hours for odd i (overlapping), [2k, 2k+1) for even i
        for k in range(i):
            st = node("POD", kind="Pod", id=f"s{i}-{k}")
            lo, hi = (k, k + 2) if i % 2 else (2 * k, 2 * k + 1)
            rel("HasState", p, st, tmin=ms_to_ts(T0 + lo * H), tmax=ms_to_ts(T0 + hi * H))
    for j, h in enumerate(hosts):                     # Node 0: five disjoint intervals; Node 1: none
        for k in range(5 if j == 0 else 0):
            st = node("NODE", kind="Node", id=f"ns{k}")
            rel("HasState", h, st, tmin=ms_to_ts(T0 + 3 * k * H), tmax=ms_to_ts(T0 + (3 * k + 2) * H))
    return lines


@pytest.mark.parametrize("family", ["fused", "split"])
def test_states_of_entities_with_up_to_seven_intervals(csr_oracle_lib, family):
    from k8s_llm_rca_b200 import engine, ingest
    snap = ingest.load_dump(_dump())
    orc = csr_oracle_lib.CsrOracle(snap)
    eng = engine.Engine(snap)
    eng.set_fused_threshold(1 << 40 if family == "fused" else 0)
    K = snap.kind_index
    pods = snap.nodes_of_kind(K("Pod")).astype(np.int32)
    seen_cols = set()
    for half_hours in range(-2, 34):                  # every half hour from before the first interval to after the last
        t = T0 + half_hours * H // 2
        for t_ in (t, t - 1):                         # ... and the millisecond before (the half-open end)
            ts = np.full(len(pods), t_, np.int64)
            for mode in (True, "mask"):
                res = eng.expand(engine.varlen_plan("out", 1, dest_kind=K("Node"), with_states=mode), pods, ts)
                got = sort_rows(res.rows(1))
                want, _, _ = orc.varlen_step(pods, ts, "out", 1, dest_kind=K("Node"))
                want = sort_rows(want[1])
                if mode is True:
                    assert np.array_equal(got, want), (half_hours, t_)
                    seen_cols.update(got[:, 3].tolist())
                else:
                    assert np.array_equal(got[:, :3], want[:, :3]) and np.array_equal(got[:, 3], want[:, 5])
    assert -1 in seen_cols and len(seen_cols) > 15    # states from the first, the second/third and the later intervals were returned
    eng.close()

"""Dump -> snapshot -> dump round trip, and the loaded snapshot answers queries like the original."""
import numpy as np

from oracle_backend import oracle_executor, record_key


def test_round_trip_reproduces_the_arrays(snap_small):
    from k8s_llm_rca_b200 import ingest
    lines = ingest.dump_snapshot(snap_small)
    loaded = ingest.load_dump(lines)
    assert [k.label for k in loaded.kinds] == [k.label for k in snap_small.kinds]
    assert [(e.rel_type, e.src_kind, e.dst_kind, e.key) for e in loaded.etypes] == \
           [(e.rel_type, e.src_kind, e.dst_kind, e.key) for e in snap_small.etypes]
    for a in ("kind_start", "out_rng", "out_dst", "in_rng", "in_adj", "und_rng", "und_adj", "typed_out", "typed_in",
              "st_off", "st_tmin", "st_tmax", "st_node", "rel_et"):
        assert np.array_equal(getattr(loaded, a), getattr(snap_small, a)), a
    assert ingest.dump_snapshot(loaded) == lines


def test_loaded_snapshot_answers_like_the_original(snap_small, tmp_path):
    from k8s_llm_rca_b200 import ingest
    from test_cypher import S2, S3
    path = str(tmp_path / "dump.jsonl")
    ingest.dump_snapshot(snap_small, path)
    loaded = ingest.load_dump(path)
    for a, b in ((oracle_executor(snap_small), oracle_executor(loaded)),
                 (oracle_executor(snap_small.derive_metagraph()), oracle_executor(loaded.derive_metagraph()))):
        for q, par in ((S3.format(arrow="-"), {"srcKind": "Pod", "destKind": "nfs", "intermediateKinds": []}),
                       (S2, {"message": snap_small.props.msgs[2]}),
                       ("MATCH (n1) WHERE n1.category IN ['NativeEntity'] RETURN n1.kind AS kind", None)):
            ra, rb = a.run_query(q, par), b.run_query(q, par)
            assert sorted(map(record_key, ra)) == sorted(map(record_key, rb))
    # the metagraph of the loaded graph has the derived categories
    kinds = {k.name: k.category for k in loaded.derive_metagraph().kinds}
    assert kinds["Pod"] == "NativeEntity" and kinds["nfs"] == "ExternalEntity"


def _same_arrays(a, b):
    for name in ("kind_start", "out_rng", "out_dst", "in_rng", "in_adj", "und_rng", "und_adj", "typed_out", "typed_in",
                 "st_off", "st_tmin", "st_tmax", "st_node", "rel_et"):
        assert np.array_equal(getattr(a, name), getattr(b, name)), name


def test_csv_round_trip_in_the_import_tool_style(snap_small, tmp_path):
    """nodes.csv / relationships.csv with `:ID,:LABEL` / `:START_ID,:END_ID,:TYPE` headers -> the same snapshot."""
    from k8s_llm_rca_b200 import ingest
    npath, rpath = ingest.dump_csv(snap_small, str(tmp_path / "csv"))
    loaded = ingest.load_csv(npath, rpath)
    _same_arrays(loaded, snap_small)
    assert [k.label for k in loaded.kinds] == [k.label for k in snap_small.kinds]
    import json                                  # same records; a CSV row has the columns' property order, so compare parsed
    assert [json.loads(x) for x in ingest.dump_snapshot(loaded)] == [json.loads(x) for x in ingest.dump_snapshot(snap_small)]
    # a directory of CSV files is a dump too, and so is the uri naming it
    from k8s_llm_rca_b200 import executor
    again = executor.resolve_snapshot("dump://" + str(tmp_path / "csv"))
    _same_arrays(again, snap_small)
    meta = executor.resolve_snapshot("dump://" + str(tmp_path / "csv") + "#metagraph")
    assert {k.name for k in meta.kinds} >= {"Pod", "nfs", "Node"}


def test_csv_headers_split_files_types_and_the_single_file_style(tmp_path):
    """Several node files, id spaces, typed and ignored columns, several labels; and the one-file export style
    (`_id,_labels,...,_start,_end,_type`).  Both describe the same little graph as the JSON-lines dump."""
    from k8s_llm_rca_b200 import ingest
    d = tmp_path / "split"
    d.mkdir()
    (d / "nodes_pod.csv").write_text("uid:ID(Obj),:LABEL,kind2,isNative,name2,replicas:int,note:IGNORE\n"
                                     "p1,Pod,Pod,true,pod-1,3,x\n"
                                     "p2,Pod;Workload,Pod,true,\"pod, two\",,y\n")
    (d / "nodes_rest.csv").write_text(":ID(Obj),:LABEL,kind2,isNative,name2,tag,path,kind,id\n"
                                      "n1,Node,Node,true,node-1,,,,\n"
                                      "x1,nfs,,false,,nfs,/exports/a,,\n"
                                      "s1,POD,,,,,,Pod,p1\n")
    (d / "rels.csv").write_text(":START_ID(Obj),:END_ID(Obj),:TYPE,key,tmin,tmax\n"
                                "p1,n1,ReferInternal,spec_nodeName,,\n"
                                "p2,n1,ReferInternal,spec_nodeName,,\n"
                                "p2,x1,UseExternal,spec_nfs_path,,\n"
                                "p1,s1,HasState,,2024-01-01 00:00:00.000,2024-01-02 00:00:00.000\n")
    a = ingest.load_path(str(d))
    assert a.n_nodes == 5 and a.n_rels == 4
    pod, node = a.kind_index("Pod"), a.kind_index("Node")
    lo, hi = a.kind_range(pod)
    props = [a.props.node_props(v) for v in range(lo, hi)]
    assert [p["name2"] for p in props] == ["pod-1", "pod, two"]
    assert props[0]["replicas"] == 3 and "replicas" not in props[1] and "note" not in props[0]
    assert props[0]["uid"] == "p1"                                   # a named id column stays a property
    assert list(a.props.node_labels(lo + 1)) == ["Pod", "Workload"]
    assert len(a.find_etypes(rel_type="ReferInternal", src_kind=pod, dst_kind=node, key="spec_nodeName")) == 1
    assert a.st_off[lo + 1] - a.st_off[lo] == 1                      # p1 has one HasState interval, p2 none
    assert a.kinds[a.kind_index("nfs")].category == "ExternalEntity"

    one = tmp_path / "all.csv"
    one.write_text("_id,_labels,kind2,isNative,name2,replicas:int,tag,path,kind,id,_start,_end,_type,key,tmin,tmax\n"
                   "p1,:Pod,Pod,true,pod-1,3,,,,,,,,,,\n"
                   "p2,:Pod:Workload,Pod,true,\"pod, two\",,,,,,,,,,,\n"
                   "n1,:Node,Node,true,node-1,,,,,,,,,,,\n"
                   "x1,:nfs,,false,,,nfs,/exports/a,,,,,,,,\n"
                   "s1,:POD,,,,,,,Pod,p1,,,,,,\n"
                   ",,,,,,,,,,p1,n1,ReferInternal,spec_nodeName,,\n"
                   ",,,,,,,,,,p2,n1,ReferInternal,spec_nodeName,,\n"
                   ",,,,,,,,,,p2,x1,UseExternal,spec_nfs_path,,\n"
                   ",,,,,,,,,,p1,s1,HasState,,2024-01-01 00:00:00.000,2024-01-02 00:00:00.000\n")
    b = ingest.load_path(str(one))
    _same_arrays(a, b)
    drop_uid = [{k: v for k, v in a.props.node_props(v).items() if k != "uid"} for v in range(a.n_nodes)]
    assert drop_uid == [dict(b.props.node_props(v)) for v in range(b.n_nodes)]


def test_a_dump_that_is_not_there_is_a_connectivity_failure(tmp_path):
    import pytest
    from k8s_llm_rca_b200 import executor
    with pytest.raises(executor.ServiceUnavailable):
        executor.resolve_snapshot("dump://" + str(tmp_path / "nothing-here.jsonl"))
    with pytest.raises(executor.ServiceUnavailable):
        executor.resolve_snapshot("dump://" + str(tmp_path))         # a directory without CSV files

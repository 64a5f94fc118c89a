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

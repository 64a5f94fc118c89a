"""Host-side mirror of the reference's module tree, for boxes where /root/reference and the neo4j
driver do not exist.

    from k8s_llm_rca_b200 import dropin
    dropin.install()            # puts this directory on sys.path
    from common.neo4j_query_executor import Neo4jQueryExecutor           # same names as the reference
    from find_metapath.find_srckind_metapath_neo4j import find_srcKind, find_metapath
    from generate_query.generate_query import human_generate_cypher_query, run_and_filter_query
    from check_state.analyze_root_cause import check_statepath
    import neo4j                # the type shim (GraphDatabase, graph.Node, exceptions.CypherSyntaxError)

`install(shim_only=True)` adds only the `neo4j` shim, so that the REFERENCE's own modules
(common/neo4j_query_executor.py, check_state/analyze_root_cause.py:3) import and run unmodified on
top of the GPU executor.
"""
import importlib.util
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))


def install(shim_only: bool = False, force_shim: bool = False) -> None:
    have_real = importlib.util.find_spec("neo4j") is not None and "k8s-llm-rca_b200" not in (
        getattr(importlib.util.find_spec("neo4j"), "origin", "") or "")
    if shim_only:
        if not have_real or force_shim:
            _register_shim()
        return
    if HERE not in sys.path:
        sys.path.insert(0, HERE)
    if have_real and not force_shim:
        return
    _register_shim()


def _register_shim():
    import importlib
    cur = sys.modules.get("neo4j")
    if cur is not None and "k8s-llm-rca_b200" in (getattr(cur, "__file__", "") or "") and "neo4j.graph" in sys.modules:
        return                                           # already in place: nothing to load again
    for name in ("neo4j", "neo4j.graph", "neo4j.exceptions"):
        if name in sys.modules and "k8s-llm-rca_b200" not in (getattr(sys.modules[name], "__file__", "") or ""):
            del sys.modules[name]
    spec = importlib.util.spec_from_file_location("neo4j", os.path.join(HERE, "neo4j", "__init__.py"),
                                                  submodule_search_locations=[os.path.join(HERE, "neo4j")])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["neo4j"] = mod
    spec.loader.exec_module(mod)

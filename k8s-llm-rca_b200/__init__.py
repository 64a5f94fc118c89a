"""B200-native path engine for the graph hot path of freiris/k8s-llm-rca.

Sub-modules: snapshot (CSR/typed-edge data format), synth (seeded generator), props (property
stores), engine (ctypes binding of the C-ABI library in csrc/), cypher (Cypher-subset front end),
executor (drop-in for the reference's Neo4jQueryExecutor), dropin/ (host-side mirror of the
reference's stage functions and a `neo4j`-shaped type shim).
"""
__version__ = "0.1.0"

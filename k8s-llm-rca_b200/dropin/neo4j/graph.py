"""neo4j.graph: the entity classes results are made of."""
from k8s_llm_rca_b200.graphtypes import Entity, Node, Path, Relationship  # noqa: F401

"""neo4j.exceptions: the two classes the reference's drivers catch or may see."""
from k8s_llm_rca_b200.cypher import CypherSyntaxError  # noqa: F401
from k8s_llm_rca_b200.executor import ServiceUnavailable  # noqa: F401


class Neo4jError(Exception):
    pass


ClientError = CypherSyntaxError

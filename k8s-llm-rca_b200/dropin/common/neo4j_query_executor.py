"""Mirror of reference common/neo4j_query_executor.py:6-24 - same class name, constructor and
methods, so `from common.neo4j_query_executor import Neo4jQueryExecutor` keeps working.  There is
no bolt driver underneath: the executor IS the GPU executor; `uri` names a snapshot."""
from k8s_llm_rca_b200.executor import GpuQueryExecutor


class Neo4jQueryExecutor(GpuQueryExecutor):
    def __init__(self, uri, user=None, password=None):
        super().__init__(uri, user, password)

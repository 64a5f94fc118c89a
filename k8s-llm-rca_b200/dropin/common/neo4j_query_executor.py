"""Mirror of reference common/neo4j_query_executor.py:6-24 - same class name, constructor and
methods, so `from common.neo4j_query_executor import Neo4jQueryExecutor` keeps working.

The executor goes through whatever module is importable as `neo4j`: on a box without the real driver
that is this repository's shim (dropin/neo4j), whose `GraphDatabase.driver(uri)` opens the GPU executor on
the snapshot the uri names - or is an alias of (executor.register_uri_alias), which is how the bolt
addresses hard-wired in the reference's drivers (test_all.py:21-22) keep working."""
import neo4j


class Neo4jQueryExecutor:
    def __init__(self, uri, user=None, password=None):
        self._link = neo4j.GraphDatabase.driver(uri, auth=(user, password))
        self._link.verify_connectivity()               # an unknown uri fails here, like an unreachable server

    @property
    def snapshot(self):
        """The Snapshot behind the shim's executor (None with a real driver)."""
        ex = getattr(self._link, "_ex", None)
        return getattr(ex, "snapshot", None)

    def run_query(self, query, parameters=None):
        session = self._link.session()
        try:
            return [row for row in session.run(query, parameters)]      # fully materialised, as the reference's list(result)
        finally:
            session.close()

    def close(self):
        link, self._link = self._link, None
        if link is not None:
            link.close()

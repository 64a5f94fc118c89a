"""`neo4j`-shaped module backed by the GPU executor.

Only what the reference touches: `GraphDatabase.driver(uri, auth=...)` with `verify_connectivity()`,
`session()` as a context manager with `run(query, parameters)`, `close()`
(common/neo4j_query_executor.py:7-21); `neo4j.graph.Node` for isinstance tests
(check_state/analyze_root_cause.py:85,97); `neo4j.exceptions.CypherSyntaxError` (test_all.py:109).
The uri is a snapshot uri (see k8s-llm-rca_b200/executor.py), not a bolt address.
"""
from k8s_llm_rca_b200.executor import GpuQueryExecutor as _Executor
from k8s_llm_rca_b200.graphtypes import Record  # noqa: F401

from . import exceptions, graph  # noqa: F401

__version__ = "0.0-rca-b200-shim"

_factory = _Executor


def set_executor_factory(factory=None):
    """What `GraphDatabase.driver(uri)` opens: `factory(uri, user, password)` -> an object with run_query / close.
    Default (and what None restores): the GPU executor."""
    global _factory
    _factory = factory or _Executor


class Result(list):
    """The driver's Result is a cursor; the reference immediately does list(result)."""

    def consume(self):
        return None

    def single(self):
        return self[0] if self else None


class Session:
    def __init__(self, executor):
        self._ex = executor

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False

    def run(self, query, parameters=None, **kw):
        p = dict(parameters or {})
        p.update(kw)
        return Result(self._ex.run_query(query, p))

    def close(self):
        pass


class Driver:
    def __init__(self, uri, auth=None, **kw):
        self._uri, self._kw, self._ex = uri, kw, None

    def _executor(self):
        if self._ex is None:
            try:
                self._ex = _factory(self._uri, None, None)
            except Exception as exc:
                raise exceptions.ServiceUnavailable(str(exc)) from exc
        return self._ex

    def verify_connectivity(self, **kw):
        self._executor()

    def session(self, **kw):
        return Session(self._executor())

    def close(self):
        if self._ex is not None:
            self._ex.close()
            self._ex = None


class GraphDatabase:
    @staticmethod
    def driver(uri, auth=None, **kw):
        return Driver(uri, auth=auth, **kw)

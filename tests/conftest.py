import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def sort_rows(a):
    a = np.asarray(a)
    if a.shape[0] == 0:
        return a
    return a[np.lexsort(a.T[::-1])]


def rows_to_set(levels):
    """{length: rows} -> sorted list of (nodes tuple, rels tuple), the canonical result form."""
    out = []
    for s, rows in levels.items():
        for r in np.asarray(rows)[:, :2 * s + 1].tolist():
            out.append((tuple(r[:s + 1]), tuple(r[s + 1:])))
    return sorted(out)


@pytest.fixture(scope="session")
def snap_small():
    from k8s_llm_rca_b200 import synth
    return synth.generate(300, 30, seed=7)


@pytest.fixture(scope="session")
def snap_1k():
    from k8s_llm_rca_b200 import synth
    return synth.generate(1000, 100, seed=synth.config_seed(0))


@pytest.fixture(scope="session")
def snap_10k():
    from k8s_llm_rca_b200 import synth
    return synth.generate(10_000, 1000, seed=synth.config_seed(1))


@pytest.fixture(scope="session")
def csr_oracle_lib():
    from oracle import csr
    csr.build()
    return csr

"""Multi-GPU host logic on CPU: seed partitioning, and a world_size-2 gloo job that runs the
collectives of k8s_llm_rca_b200.parallel for real (all-gather of counts, gather of rows)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT


def test_partition_by_work_is_contiguous_and_balanced():
    from k8s_llm_rca_b200.parallel import partition_by_work
    rng = np.random.default_rng(0)
    for n, world in ((0, 4), (3, 8), (1000, 1), (1000, 2), (1000, 8), (100_000, 8)):
        w = rng.integers(0, 50, size=n).astype(float)
        if n > 10:
            w[rng.integers(0, n, size=3)] = 5000.0          # hubs
        parts = partition_by_work(w, world)
        assert len(parts) == world and parts[0][0] == 0 and parts[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(parts, parts[1:])) and all(lo <= hi for lo, hi in parts)
        if n >= 1000 and world > 1:
            sums = np.array([w[lo:hi].sum() for lo, hi in parts])
            assert sums.max() <= w.sum() / world + w.max() + 1


def test_first_hop_degree_matches_snapshot(snap_small):
    from k8s_llm_rca_b200.parallel import first_hop_degree
    s = snap_small
    seeds = np.array([0, 5, s.n_nodes - 1, -3, s.n_nodes + 9])
    d = first_hop_degree(s, seeds, "und", True)
    assert d[0] == s.und_rng[0, 1] - s.und_rng[0, 0] and d[3] == 0 and d[4] == 0
    assert first_hop_degree(s, seeds[:3], "out", False)[1] == s.out_rng[6, 0] - s.out_rng[5, 0]


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_expansion_over_gloo(world):
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="2")
    port = 29500 + (os.getpid() % 2000) + world
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "dist_worker.py")]
    p = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-4000:]
    assert p.stdout.count(" ok") == world

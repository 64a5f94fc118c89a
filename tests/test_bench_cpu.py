"""bench.py on a box without a GPU: the reference arm (--impl reference) prints the contract's JSON line from the C
restatement, the roofline's DRAM figure is quoted only from a capture of the current kernel sources, and the b200 arm
refuses to run without a device instead of falling back to anything."""
import json
import os
import subprocess
import sys

import pytest

from conftest import ROOT


def _run(*args, timeout=240):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, timeout=timeout)


def test_reference_arm_prints_the_contract_line():
    p = _run("--impl", "reference", "--objects", "3000", "--events", "400", "--steps", "2", "--warmup", "1")
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [ln for ln in p.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1                                    # ONE JSON line on stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "path-instances/s" and d["unit"] == "instances/s"
    assert d["higher_is_better"] is True and d["steps"] == 2 and d["value"] > 0
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["pruned"]["value"] > 0
    assert d["e2e"] == {"value": d["value"], "unit": "instances/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["incidents_per_step"] > 0 and "config3" in d["config"]["workload"]


def test_reference_arm_other_ranks_do_nothing():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                       capture_output=True, text=True, timeout=120, env=env)
    assert p.returncode == 0 and p.stdout.strip() == ""


def test_b200_arm_has_no_cpu_path():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    p = _run("--steps", "1", "--warmup", "1", timeout=120)
    assert p.returncode != 0 and "no CUDA device" in (p.stderr + p.stdout)


def test_traffic_is_quoted_only_from_the_current_sources(tmp_path, monkeypatch):
    sys.path.insert(0, ROOT)
    import bench
    got, note = bench.captured_traffic("varlen/level3")
    stamp = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    if stamp["csrc_sha"] == bench.csrc_digest():
        assert got == stamp["varlen/level3"] and stamp["commit"] in note
    else:
        assert got is None and "other kernel sources" in note
    monkeypatch.setattr(bench, "csrc_digest", lambda: "0" * 16)      # sources 'changed': the capture must not be quoted
    got, note = bench.captured_traffic("varlen/level3")
    assert got is None and "other kernel sources" in note

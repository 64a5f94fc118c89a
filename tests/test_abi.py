"""The C-ABI library loads and exports every symbol include/rca_b200.h declares (no compute)."""
import ctypes
import os
import re

from conftest import ROOT


def declared_functions():
    text = open(os.path.join(ROOT, "include", "rca_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rca_[a-z_0-9]+)\s*\(", text)))


def test_header_declares_the_boundary():
    names = declared_functions()
    for must in ("rca_open", "rca_close", "rca_expand", "rca_result_copy", "rca_state_lookup", "rca_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from k8s_llm_rca_b200 import engine
    lib = engine.load_library()
    for name in declared_functions():
        assert hasattr(lib, name), f"{name} is declared in include/rca_b200.h but not exported"
    assert b"sm_100a" in lib.rca_version()
    assert set(engine.EXPORTS) == set(declared_functions())


def test_struct_layouts_match_header():
    """ctypes mirrors must have the sizes the C compiler gives the header's structs."""
    import subprocess
    import tempfile
    from k8s_llm_rca_b200 import engine
    src = '#include <stdio.h>\n#include "rca_b200.h"\nint main(){printf("%zu %zu %zu\\n", sizeof(rca_snapshot_t), sizeof(rca_plan_t), sizeof(rca_stats_t));return 0;}\n'
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "s.c")
        open(c, "w").write(src)
        exe = os.path.join(d, "s")
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), c, "-o", exe])
        a, b, s = map(int, subprocess.check_output([exe]).split())
    assert ctypes.sizeof(engine.SnapshotDesc) == a
    assert ctypes.sizeof(engine.PlanDesc) == b
    assert ctypes.sizeof(engine.StatsDesc) == s


def test_open_without_gpu_fails_loudly():
    """No CPU fallback: on a box without a B200 the engine must refuse, not emulate."""
    import pytest
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        pytest.skip("a GPU is present")
    from k8s_llm_rca_b200 import engine, synth
    snap = synth.generate(200, 10, seed=1)
    with pytest.raises(engine.RcaError) as ei:
        engine.Engine(snap)
    assert "no CUDA device" in str(ei.value) or "CPU path" in str(ei.value)

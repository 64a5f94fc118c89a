"""ctypes binding of the C-ABI library (include/rca_b200.h, csrc/rca_b200.cu).

This is the only door from Python to the kernels.  There is no CPU implementation behind it:
if the shared library is missing, or no B200 is visible, the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

MAX_HOPS = 8
MAX_KINDS = 256
STATE_LIMIT = 10
DIR_OUT, DIR_IN, DIR_UND = 0, 1, 2
PLAN_VARLEN, PLAN_CHAIN = 0, 1
_DIRS = {"out": DIR_OUT, "in": DIR_IN, "und": DIR_UND}

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "librca_b200.so")

EXPORTS = ("rca_open", "rca_close", "rca_last_error", "rca_set_stream", "rca_set_fused_threshold",
           "rca_set_chain_walk_fanout", "rca_set_unit_steps", "rca_set_wide_entries", "rca_set_graphs", "rca_graph_bytes", "rca_reserve",
           "rca_expand", "rca_expand_many", "rca_result_select", "rca_expand_static",
           "rca_result_level", "rca_result_copy", "rca_set_double_buffer", "rca_result_copy_async", "rca_result_wait", "rca_state_lookup", "rca_state_window", "rca_state_window_kind", "rca_text_load",
           "rca_text_contains", "rca_text_match_pairs", "rca_peer_handle", "rca_peer_connect", "rca_peer_allgather",
           "rca_peer_wait", "rca_peer_set_timeout", "rca_gather_window", "rca_gather_connect", "rca_gather_rows",
           "rca_version")


class RcaError(RuntimeError):
    def __init__(self, code, text):
        super().__init__(f"rca_b200 error {code}: {text}")
        self.code = code


class SnapshotDesc(C.Structure):
    _fields_ = [("n_nodes", C.c_int32), ("n_kinds", C.c_int32), ("n_etypes", C.c_int32), ("reserved", C.c_int32),
                ("n_rels", C.c_int64), ("n_und", C.c_int64), ("n_states", C.c_int64),
                ("n_typed_out", C.c_int64), ("n_typed_in", C.c_int64),
                ("kind_start", C.c_void_p), ("et_src", C.c_void_p), ("et_dst", C.c_void_p), ("et_cls", C.c_void_p),
                ("out_rng", C.c_void_p), ("out_dst", C.c_void_p), ("in_rng", C.c_void_p), ("in_adj", C.c_void_p),
                ("und_rng", C.c_void_p), ("und_adj", C.c_void_p),
                ("typed_out", C.c_void_p), ("typed_out_base", C.c_void_p),
                ("typed_in", C.c_void_p), ("typed_in_base", C.c_void_p),
                ("st_off", C.c_void_p), ("st_tmin", C.c_void_p), ("st_tmax", C.c_void_p), ("st_node", C.c_void_p),
                ("rel_et", C.c_void_p)]


class PlanDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("n_steps", C.c_int32),
                ("step_etype", C.c_int32 * MAX_HOPS), ("step_dir", C.c_int32 * MAX_HOPS),
                ("direction", C.c_int32), ("min_hops", C.c_int32), ("max_hops", C.c_int32),
                ("entity_only", C.c_int32), ("node_simple", C.c_int32), ("dest_kind", C.c_int32),
                ("has_inter", C.c_int32), ("prune", C.c_int32),
                ("pos_kind", C.c_int32 * (MAX_HOPS + 1)),
                ("kind_excl", C.c_uint8 * MAX_KINDS), ("kind_inter", C.c_uint8 * MAX_KINDS),
                ("with_states", C.c_int32), ("count_only", C.c_int32), ("shortest_only", C.c_int32)]


class StatsDesc(C.Structure):
    _fields_ = [("frontier", C.c_int64 * (MAX_HOPS + 1)), ("scanned", C.c_int64 * (MAX_HOPS + 1)),
                ("emitted", C.c_int64 * (MAX_HOPS + 1)), ("row_len", C.c_int32 * (MAX_HOPS + 1)),
                ("total", C.c_int64), ("kernel_ms", C.c_double), ("level_ms", C.c_double * (MAX_HOPS + 1)), ("launches", C.c_int32), ("retries", C.c_int32)]


_lib = None


def load_library(path: str | None = None):
    """Load librca_b200.so.  Raises if it has not been built (python -c 'import __graft_entry__ as
    g; g.build()') - there is deliberately nothing to fall back to."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or os.environ.get("RCA_B200_LIB") or LIB_PATH     # the env var is for kernel experiments only
    if not os.path.exists(p):
        raise RuntimeError(f"{p} is missing: build the CUDA library first (__graft_entry__.build()); "
                           "this package has no CPU path")
    lib = C.CDLL(p)
    lib.rca_last_error.restype = C.c_char_p
    lib.rca_last_error.argtypes = [C.c_void_p]
    lib.rca_version.restype = C.c_char_p
    lib.rca_open.argtypes = [C.POINTER(SnapshotDesc), C.c_int, C.POINTER(C.c_void_p)]
    lib.rca_close.argtypes = [C.c_void_p]
    lib.rca_set_stream.argtypes = [C.c_void_p, C.c_uint64]
    lib.rca_reserve.argtypes = [C.c_void_p, C.c_int64, C.c_int64]
    lib.rca_set_fused_threshold.argtypes = [C.c_void_p, C.c_int64]
    lib.rca_set_chain_walk_fanout.argtypes = [C.c_void_p, C.c_int64]
    lib.rca_set_unit_steps.argtypes = [C.c_void_p, C.c_int]
    lib.rca_set_wide_entries.argtypes = [C.c_void_p, C.c_int]
    lib.rca_set_graphs.argtypes = [C.c_void_p, C.c_int]
    lib.rca_graph_bytes.argtypes = [C.c_void_p]
    lib.rca_graph_bytes.restype = C.c_int64
    lib.rca_expand.argtypes = [C.c_void_p, C.POINTER(PlanDesc), C.c_void_p, C.c_int64, C.c_void_p, C.c_int,
                               C.POINTER(StatsDesc)]
    lib.rca_expand_static.argtypes = [C.c_void_p, C.POINTER(PlanDesc), C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p,
                                      C.c_int64, C.POINTER(StatsDesc)]
    lib.rca_expand_many.argtypes = [C.c_void_p, C.POINTER(PlanDesc), C.c_int32, C.POINTER(C.c_void_p), C.POINTER(C.c_int64),
                                    C.POINTER(C.c_void_p), C.c_int, C.POINTER(StatsDesc)]
    lib.rca_result_select.argtypes = [C.c_void_p, C.c_int32]
    lib.rca_result_level.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_int64),
                                     C.POINTER(C.c_int32)]
    lib.rca_result_copy.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    lib.rca_state_lookup.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p,
                                     C.c_void_p]
    lib.rca_state_window.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p,
                                     C.c_void_p, C.c_void_p]
    lib.rca_state_window_kind.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int32,
                                          C.c_void_p, C.c_void_p, C.c_void_p]
    lib.rca_text_load.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]
    lib.rca_set_double_buffer.argtypes = [C.c_void_p, C.c_int]
    lib.rca_result_copy_async.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    lib.rca_result_wait.argtypes = [C.c_void_p]
    lib.rca_peer_handle.argtypes = [C.c_void_p, C.c_int32, C.c_void_p]
    lib.rca_peer_connect.argtypes = [C.c_void_p, C.c_int32, C.c_void_p]
    lib.rca_peer_allgather.argtypes = [C.c_void_p, C.c_int64, C.c_void_p]
    lib.rca_peer_wait.argtypes = [C.c_void_p, C.c_void_p]
    lib.rca_peer_set_timeout.argtypes = [C.c_void_p, C.c_int64]
    lib.rca_gather_window.argtypes = [C.c_void_p, C.c_int64, C.c_void_p]
    lib.rca_gather_connect.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64]
    lib.rca_gather_rows.argtypes = [C.c_void_p, C.c_int, C.c_int64, C.POINTER(C.c_void_p), C.POINTER(C.c_int64), C.c_void_p,
                                    C.POINTER(C.c_double)]
    lib.rca_text_match_pairs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_int64, C.c_void_p]
    lib.rca_text_contains.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64,
                                      C.POINTER(C.c_int64)]
    if path is None:
        _lib = lib
    return lib


def varlen_plan(direction, max_hops, min_hops=1, dest_kind=-1, excl_kinds=(), inter_kinds=(), entity_only=False,
                node_simple=True, pos_kind=None, prune=True, with_states=False, count_only=False,
                shortest_only=False) -> PlanDesc:
    """Variable-length path plan - the reference's metapath statements,
    find_metapath/find_srckind_metapath_neo4j.py:95-129."""
    p = PlanDesc()
    p.kind = PLAN_VARLEN
    p.direction = _DIRS[direction] if isinstance(direction, str) else int(direction)
    p.min_hops, p.max_hops = int(min_hops), int(max_hops)
    p.entity_only, p.node_simple = int(bool(entity_only)), int(bool(node_simple))
    p.dest_kind = int(dest_kind)
    p.has_inter = int(len(inter_kinds) > 0)
    p.prune = int(bool(prune))
    for i in range(MAX_HOPS + 1):
        p.pos_kind[i] = -1
    if pos_kind is not None:
        for i, k in enumerate(pos_kind):
            p.pos_kind[i] = int(k)
    for k in excl_kinds:
        p.kind_excl[int(k)] = 1
    for k in inter_kinds:
        p.kind_inter[int(k)] = 1
    p.with_states = _states_mode(with_states)
    p.count_only = int(bool(count_only))
    p.shortest_only = int(bool(shortest_only))
    return p


def chain_plan(etypes, dirs, with_states=False, count_only=False) -> PlanDesc:
    """Typed chain plan - one step per MATCH clause of the instance query,
    generate_query/generate_query.py:241-248."""
    assert len(etypes) == len(dirs) and 1 <= len(etypes) <= MAX_HOPS
    p = PlanDesc()
    p.kind = PLAN_CHAIN
    p.n_steps = len(etypes)
    for i, (t, d) in enumerate(zip(etypes, dirs)):
        p.step_etype[i] = int(t)
        p.step_dir[i] = _DIRS[d] if isinstance(d, str) else int(d)
    for i in range(MAX_HOPS + 1):
        p.pos_kind[i] = -1
    p.dest_kind = -1
    p.with_states = _states_mode(with_states)
    p.count_only = int(bool(count_only))
    return p


STATES_NONE, STATES_FULL, STATES_MASK = 0, 1, 2


def _states_mode(with_states) -> int:
    """False / True / "mask" (or the RCA_STATES_* numbers): no state columns, a state id per node plus the missing
    mask, the mask alone."""
    if with_states in ("mask", STATES_MASK) and with_states is not True:
        return STATES_MASK
    return STATES_FULL if with_states else STATES_NONE


def _dev_ptr(x):
    """(pointer, on_device, keepalive) for a numpy array, a torch tensor or a raw device pointer."""
    if x is None:
        return None, 0, None
    if isinstance(x, np.ndarray):
        a = np.ascontiguousarray(x)
        return a.ctypes.data, 0, a
    if hasattr(x, "data_ptr"):                       # torch tensor
        t = x.contiguous()
        return t.data_ptr(), int(t.is_cuda), t
    raise TypeError(f"unsupported buffer type {type(x)}")


class ExpandResult:
    """Counters of one expansion plus access to its rows.  Rows live in the engine's output arena
    until the next expand() on the same engine; rows() copies them out."""

    _ARRAYS = ("frontier", "scanned", "emitted", "row_len", "level_ms")

    def __init__(self, engine, st: StatsDesc, levels: int):
        self.engine = engine
        self.levels = levels
        self._st = st                  # the per-level arrays become lists when first asked for
        self.total = st.total
        self.kernel_ms = st.kernel_ms
        self.launches = st.launches
        self.retries = st.retries
        self._gen = engine._gen
        self.count_only = False
        self.index = 0                 # position in the submission (expand_many)

    def __getattr__(self, name):
        if name in ExpandResult._ARRAYS:
            v = list(getattr(self._st, name))[:self.levels + 1]
            setattr(self, name, v)
            return v
        raise AttributeError(name)

    def rows(self, level: int, out: np.ndarray | None = None) -> np.ndarray:
        """Copy the rows of one level to the host.  `out`: optional pre-allocated (ideally pinned)
        int32 buffer with room for n*row_len words; a view of it is returned."""
        assert self._gen == self.engine._gen, "result was overwritten by a later expand()"
        n, w = self.emitted[level], self.row_len[level]
        if self.count_only and n:
            raise ValueError("the plan ran with count_only: there are no rows to copy")
        self.engine._ck(self.engine.lib.rca_result_select(self.engine.h, self.index))
        if out is not None:
            flat = out.reshape(-1)
            assert flat.dtype == np.int32 and flat.size >= n * w and flat.flags.c_contiguous
            out = flat[:n * w].reshape(n, w)
        else:
            out = np.empty((n, w), dtype=np.int32)
        if n:
            self.engine._ck(self.engine.lib.rca_result_copy(self.engine.h, level, out.ctypes.data))
        return out

    def rows_async(self, level: int, out: np.ndarray) -> np.ndarray:
        """Enqueue the copy of one level's rows into `out` (pinned int32 memory) and return at once; the data is
        there after engine.result_wait().  Needs engine.set_double_buffer(True): the next expand() then runs
        while these rows are still travelling."""
        assert self._gen == self.engine._gen, "result was overwritten by a later expand()"
        n, w = self.emitted[level], self.row_len[level]
        flat = out.reshape(-1)
        assert flat.dtype == np.int32 and flat.size >= n * w and flat.flags.c_contiguous
        view = flat[:n * w].reshape(n, w)
        if n:
            self.engine._ck(self.engine.lib.rca_result_select(self.engine.h, self.index))
            self.engine._ck(self.engine.lib.rca_result_copy_async(self.engine.h, level, view.ctypes.data))
        return view

    def all_rows(self) -> dict:
        return {s: self.rows(s) for s in range(self.levels + 1) if self.emitted[s]}

    def device_rows(self, level: int):
        """(device pointer, n_rows, row_len) - for consumers that stay on the GPU (NCCL gather)."""
        assert self._gen == self.engine._gen, "result was overwritten by a later expand()"
        p, n, w = C.c_void_p(), C.c_int64(), C.c_int32()
        self.engine._ck(self.engine.lib.rca_result_select(self.engine.h, self.index))
        self.engine._ck(self.engine.lib.rca_result_level(self.engine.h, level, C.byref(p), C.byref(n), C.byref(w)))
        return (p.value or 0), int(n.value), int(w.value)

    def level_bytes(self, s: int) -> int:
        """Algorithmic bytes of the launch that produced level s:
        12 F_{s-1} + 4 E_s + 4 (2s+1) N_s, plus, with the fused state filter,
        (s+1) N_s (8 + 16 I + 5) with I = 1 interval inspected per node."""
        b = 12 * self.frontier[s - 1] + 4 * self.scanned[s] + 4 * (2 * s + 1) * self.emitted[s]
        if self.row_len[s] == 3 * s + 3:
            b += (s + 1) * self.emitted[s] * (8 + 16 + 5)
        elif self.row_len[s] == 2 * s + 2:               # mask-only rows: no state id written, one mask word per row
            b += (s + 1) * self.emitted[s] * (8 + 16) + 4 * self.emitted[s]
        return b

    def algorithmic_bytes(self) -> int:
        return sum(self.level_bytes(s) for s in range(1, self.levels + 1))


class Engine:
    """One snapshot resident on one GPU."""

    def __init__(self, snap, device: int = 0):
        self.lib = load_library()
        self.snap = snap
        self.device = device
        self._gen = 0
        names = ("kind_start", "et_src", "et_dst", "et_cls", "out_rng", "out_dst", "in_rng", "in_adj", "und_rng",
                 "und_adj", "typed_out", "typed_out_base", "typed_in", "typed_in_base", "st_off", "st_tmin",
                 "st_tmax", "st_node", "rel_et")
        dt = {"typed_out_base": np.int64, "typed_in_base": np.int64, "st_tmin": np.int64, "st_tmax": np.int64}
        keep = {n: np.ascontiguousarray(getattr(snap, n), dtype=dt.get(n, np.int32)) for n in names}
        d = SnapshotDesc()
        d.n_nodes, d.n_kinds, d.n_etypes = snap.n_nodes, snap.n_kinds, snap.n_etypes
        d.n_rels, d.n_und, d.n_states = snap.n_rels, len(keep["und_adj"]), len(keep["st_node"])
        d.n_typed_out, d.n_typed_in = len(keep["typed_out"]), len(keep["typed_in"])
        for n in names:
            setattr(d, n, keep[n].ctypes.data)
        h = C.c_void_p()
        rc = self.lib.rca_open(C.byref(d), device, C.byref(h))
        if rc != 0:
            raise RcaError(rc, self.lib.rca_last_error(None).decode())
        self.h = h

    def _ck(self, rc):
        if rc != 0:
            raise RcaError(rc, self.lib.rca_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            self.lib.rca_close(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, cuda_stream: int):
        self._ck(self.lib.rca_set_stream(self.h, int(cuda_stream)))

    def set_fused_threshold(self, max_seeds: int):
        """Seed sets up to this size run the single-launch-per-hop kernels (see include/rca_b200.h)."""
        self._ck(self.lib.rca_set_fused_threshold(self.h, int(max_seeds)))

    def set_chain_walk_fanout(self, max_rows_per_seed: int):
        """Typed chains with at most this per-seed fan-out run as one depth-first launch (0: never)."""
        self._ck(self.lib.rca_set_chain_walk_fanout(self.h, int(max_rows_per_seed)))

    def set_unit_steps(self, on: bool):
        """Typed steps with at most one candidate per path run through the one-lane-per-path kernels (default on)."""
        self._ck(self.lib.rca_set_unit_steps(self.h, 1 if on else 0))

    def set_wide_entries(self, on: bool):
        """The level before a final unit level stores whole paths (48-byte wide entries) instead of parent-pointer
        entries (default off; include/rca_b200.h)."""
        self._ck(self.lib.rca_set_wide_entries(self.h, 1 if on else 0))

    def set_graphs(self, on: bool):
        """Repeated submissions of expand_many are replayed as CUDA graphs (default off; include/rca_b200.h)."""
        self._ck(self.lib.rca_set_graphs(self.h, 1 if on else 0))

    def set_double_buffer(self, on: bool):
        """Submissions alternate between two output arenas, so ExpandResult.rows_async() copies can overlap the
        next submission (include/rca_b200.h)."""
        self._ck(self.lib.rca_set_double_buffer(self.h, 1 if on else 0))

    def result_wait(self):
        """All rows_async() copies have arrived."""
        self._ck(self.lib.rca_result_wait(self.h))

    def graph_bytes(self) -> int:
        """Device memory the snapshot occupies on this engine's GPU."""
        return int(self.lib.rca_graph_bytes(self.h))

    def reserve(self, frontier_entries: int, output_words: int):
        self._ck(self.lib.rca_reserve(self.h, int(frontier_entries), int(output_words)))
        self._gen += 1

    def expand(self, plan: PlanDesc, seeds, seed_ts=None) -> ExpandResult:
        sp, on_dev, k1 = _dev_ptr(seeds if not isinstance(seeds, (list, tuple)) else np.asarray(seeds, np.int32))
        if isinstance(k1, np.ndarray) and k1.dtype != np.int32:
            k1 = np.ascontiguousarray(k1, dtype=np.int32)
            sp = k1.ctypes.data
        n = int(k1.shape[0]) if k1 is not None else 0
        tp, ts_dev, k2 = _dev_ptr(seed_ts if seed_ts is None or hasattr(seed_ts, "data_ptr")
                                  else np.ascontiguousarray(seed_ts, dtype=np.int64))
        if seed_ts is not None and ts_dev != on_dev:
            raise ValueError("seeds and seed_ts must both be host or both be device buffers")
        st = StatsDesc()
        self._gen += 1
        self._ck(self.lib.rca_expand(self.h, C.byref(plan), sp, n, tp, on_dev, C.byref(st)))
        levels = plan.max_hops if plan.kind == PLAN_VARLEN else plan.n_steps
        res = ExpandResult(self, st, levels)
        res.count_only = bool(plan.count_only)
        return res

    def expand_many(self, plans, seeds_list, ts_list=None):
        """Several plans in one submission (rca_expand_many): one synchronisation for the whole set.
        seeds_list[k] / ts_list[k]: numpy arrays, or CUDA tensors for all plans.  -> [ExpandResult]."""
        n = len(plans)
        # a submission that comes again with the very same plan and buffer objects (a benchmark loop, a service replaying
        # a prepared batch) reuses its marshalled descriptor: building it is a third of the host side of a small step
        ck = (tuple(map(id, plans)), tuple(map(id, seeds_list)), tuple(map(id, ts_list)) if ts_list is not None else None)
        hit = self._many_cache.get(ck) if hasattr(self, "_many_cache") else None
        if hit is not None and all(p == q for p, q in zip(hit["ptrs"], self._live_ptrs(seeds_list, ts_list))):
            st = (StatsDesc * n)()
            self._gen += 1
            arr = (PlanDesc * n)(*plans)             # the plans themselves are copied afresh: they may have been edited in place
            self._ck(self.lib.rca_expand_many(self.h, arr, n, hit["sp"], hit["ns"], hit["tp"], hit["dev"], st))
            return self._many_results(plans, st, [p.max_hops if p.kind == PLAN_VARLEN else p.n_steps for p in plans])
        arr = (PlanDesc * n)(*plans)
        keep, sp, tp, ns, devs = [], (C.c_void_p * n)(), (C.c_void_p * n)(), (C.c_int64 * n)(), set()
        seen = {}                       # plans of one submission usually share their seed / timestamp buffers

        def ptr_of(x, dtype):
            key = id(x)
            hit = seen.get(key)
            if hit is None:
                if isinstance(x, (list, tuple)) or (isinstance(x, np.ndarray) and x.dtype != dtype):
                    x = np.ascontiguousarray(x, dtype=dtype)
                hit = seen[key] = _dev_ptr(x)
                keep.append(hit[2])
                devs.add(hit[1])
            return hit

        for k in range(n):
            ptr, _, ka = ptr_of(seeds_list[k], np.int32)
            sp[k], ns[k] = ptr, int(ka.shape[0])
            t = ts_list[k] if ts_list is not None else None
            if t is not None:
                tp[k] = ptr_of(t, np.int64)[0]
        if len(devs) > 1:
            raise ValueError("all seed / timestamp buffers must be host arrays or all device tensors")
        st = (StatsDesc * n)()
        self._gen += 1
        dev = devs.pop() if devs else 0
        self._ck(self.lib.rca_expand_many(self.h, arr, n, sp, ns, tp, dev, st))
        levels = [p.max_hops if p.kind == PLAN_VARLEN else p.n_steps for p in plans]
        if all(hasattr(x, "data_ptr") for x in seeds_list):          # tensors only: their identity and address can be re-checked cheaply
            if not hasattr(self, "_many_cache"):
                self._many_cache = {}
            if len(self._many_cache) >= 4:
                self._many_cache.pop(next(iter(self._many_cache)))
            self._many_cache[ck] = {"sp": sp, "tp": tp, "ns": ns, "dev": dev, "levels": levels,
                                    "keep": (keep, list(plans), list(seeds_list), list(ts_list) if ts_list is not None else None),
                                    "ptrs": self._live_ptrs(seeds_list, ts_list)}
        return self._many_results(plans, st, levels)

    @staticmethod
    def _live_ptrs(seeds_list, ts_list):
        """(address, length) of every buffer as it is NOW - what a cached descriptor is re-validated against."""
        out = [(x.data_ptr(), x.numel()) if hasattr(x, "data_ptr") else None for x in seeds_list]
        if ts_list is not None:
            out += [(x.data_ptr(), x.numel()) if hasattr(x, "data_ptr") else None for x in ts_list]
        return out

    def _many_results(self, plans, st, levels):
        out = []
        for k in range(len(plans)):
            r = ExpandResult(self, st[k], levels[k])
            r.count_only = bool(plans[k].count_only)
            r.index = k
            out.append(r)
        return out

    def expand_static(self, plan: PlanDesc, dev_seeds, dev_seed_ts, dev_rows):
        """rca_expand_static: expansion that allocates nothing (fixed scratch inside the library; see
        include/rca_b200.h for the limits and the sharing rule).  All three buffers are device
        tensors owned by the caller: int32 seeds, int64 timestamps (or None), int32 row buffer.
        The kernels run on the engine's stream: call set_stream(torch.cuda.current_stream().cuda_stream)
        first, or synchronise after filling the tensors.
        Returns a StatsDesc; rows of length s start at word sum_{j<s} emitted[j]*row_len[j]."""
        import torch
        want = [("dev_seeds", dev_seeds, torch.int32), ("dev_rows", dev_rows, torch.int32)]
        if dev_seed_ts is not None:
            want.append(("dev_seed_ts", dev_seed_ts, torch.int64))
        for name, t, dt in want:
            if not (hasattr(t, "is_cuda") and t.is_cuda and t.is_contiguous()):
                raise ValueError(f"expand_static: {name} must be a contiguous device tensor")
            if t.dtype != dt:
                raise ValueError(f"expand_static: {name} must be {dt}, got {t.dtype}")
            if t.device.index != self.device:
                raise ValueError(f"expand_static: {name} lives on cuda:{t.device.index}, the engine on cuda:{self.device}")
        if dev_seed_ts is not None and dev_seed_ts.numel() < dev_seeds.numel():
            raise ValueError("expand_static: dev_seed_ts is shorter than dev_seeds")
        st = StatsDesc()
        self._ck(self.lib.rca_expand_static(
            self.h, C.byref(plan), dev_seeds.data_ptr(), int(dev_seeds.numel()),
            dev_seed_ts.data_ptr() if dev_seed_ts is not None else None,
            dev_rows.data_ptr(), int(dev_rows.numel()), C.byref(st)))
        return st

    def text_load(self, node_ids, messages):
        """Upload the message arena (node id, message string) for contains(); sorted by node id."""
        order = np.argsort(np.asarray(node_ids), kind="stable")
        ids = np.ascontiguousarray(np.asarray(node_ids, dtype=np.int32)[order])
        enc = [(messages[i] or "").encode("utf-8") for i in order.tolist()]
        off = np.zeros(len(enc) + 1, dtype=np.int64)
        np.cumsum([len(b) for b in enc], out=off[1:])
        blob = np.frombuffer(b"".join(enc), dtype=np.uint8) if off[-1] else np.zeros(1, np.uint8)
        blob = np.ascontiguousarray(blob)
        self._ck(self.lib.rca_text_load(self.h, blob.ctypes.data, off.ctypes.data, ids.ctypes.data, len(enc)))
        self._text_nodes = ids

    def contains(self, patterns, pairs_cap=None):
        """Batched `message CONTAINS pattern`: -> (first node id per pattern or -1, {pattern index: sorted
        node ids})."""
        enc = [t.encode("utf-8") for t in patterns]
        uniq = {}
        for b in enc:
            uniq.setdefault(b, len(uniq))
        ulist = list(uniq)
        off = np.zeros(len(ulist) + 1, dtype=np.int64)
        np.cumsum([len(b) for b in ulist], out=off[1:])
        blob = np.ascontiguousarray(np.frombuffer(b"".join(ulist), dtype=np.uint8)) if off[-1] else np.zeros(1, np.uint8)
        first = np.empty(len(ulist), np.int32)
        cap = int(pairs_cap or max(1024, 16 * len(ulist)))
        while True:
            pairs = np.empty((cap, 2), np.int32)
            n = C.c_int64(0)
            self._ck(self.lib.rca_text_contains(self.h, blob.ctypes.data, off.ctypes.data, len(ulist), first.ctypes.data,
                                                pairs.ctypes.data, cap, C.byref(n)))
            if n.value <= cap:
                break
            cap = int(n.value) + 16
        by_u = {}
        for u, node in sorted(set(map(tuple, pairs[:n.value].tolist()))):      # few pairs per call: plain Python beats np.unique
            by_u.setdefault(u, []).append(node)
        idx = [uniq[b] for b in enc]
        hits = {}
        for i, u in enumerate(idx):
            if len(ulist[u]) == 0:
                hits[i] = self._text_nodes.tolist()
            elif u in by_u:
                hits[i] = by_u[u]
        return first[idx], hits

    def contains_first(self, patterns):
        """Batched `message CONTAINS pattern`, the canonical 'LIMIT 1' answer only: the lowest node id whose message
        contains each pattern, or -1 (no list of all matches is built or copied)."""
        uniq = {}
        idx = np.fromiter((uniq.setdefault(t, len(uniq)) for t in patterns), dtype=np.int64, count=len(patterns))
        enc = [t.encode("utf-8") for t in uniq]
        off = np.zeros(len(enc) + 1, dtype=np.int64)
        np.cumsum(np.fromiter(map(len, enc), dtype=np.int64, count=len(enc)), out=off[1:])
        blob = np.frombuffer(b"".join(enc), dtype=np.uint8) if off[-1] else np.zeros(1, np.uint8)
        first = np.empty(len(enc), np.int32)
        n = C.c_int64(0)
        if len(enc):
            self._ck(self.lib.rca_text_contains(self.h, blob.ctypes.data, off.ctypes.data, len(enc), first.ctypes.data,
                                                None, 0, C.byref(n)))
        return first[idx]

    def message_index(self, nodes):
        """Index in the loaded message arena of the messages of `nodes` (node ids given to text_load)."""
        nodes = np.asarray(nodes, dtype=np.int32)
        idx = np.searchsorted(self._text_nodes, nodes)
        if len(nodes) and ((idx >= len(self._text_nodes)).any() or (self._text_nodes[np.minimum(idx, len(self._text_nodes) - 1)] != nodes).any()):
            raise ValueError("a node has no message in the loaded arena")
        return np.ascontiguousarray(idx, dtype=np.int32)

    def match_pairs_idx(self, msg_index, needle_a, needle_b, needle_blob, needle_off):
        """rca_text_match_pairs on prepared arrays: message indices (message_index), two needle indices per pair
        (-1: none) into an arena given as (uint8 bytes, int64 offsets).  -> bool array."""
        n = len(msg_index)
        if n == 0:
            return np.zeros(0, dtype=bool)
        mi = np.ascontiguousarray(msg_index, dtype=np.int32)
        na = np.ascontiguousarray(needle_a, dtype=np.int32)
        nb = np.ascontiguousarray(needle_b, dtype=np.int32)
        blob = np.ascontiguousarray(needle_blob, dtype=np.uint8)
        off = np.ascontiguousarray(needle_off, dtype=np.int64)
        out = np.empty(n, np.uint8)
        self._ck(self.lib.rca_text_match_pairs(self.h, blob.ctypes.data, off.ctypes.data, len(off) - 1, mi.ctypes.data,
                                               na.ctypes.data, nb.ctypes.data, n, out.ctypes.data))
        return out.astype(bool)

    # -- count exchange over peer memory (include/rca_b200.h: rca_peer_*)
    PEER_HANDLE_BYTES = 64

    def peer_handle(self, world: int) -> bytes:
        buf = C.create_string_buffer(self.PEER_HANDLE_BYTES)
        self._ck(self.lib.rca_peer_handle(self.h, int(world), buf))
        self._peer_world = int(world)
        return buf.raw

    def peer_connect(self, rank: int, handles) -> None:
        """handles: the peer_handle() bytes of every rank, in rank order."""
        blob = b"".join(handles)
        assert len(blob) == self._peer_world * self.PEER_HANDLE_BYTES
        self._ck(self.lib.rca_peer_connect(self.h, int(rank), blob))

    def peer_allgather(self, value: int, wait: bool = True):
        """Publish `value` to every rank and (wait=True) return all ranks' values; wait=False only enqueues
        the exchange on the engine's stream - collect with peer_wait()."""
        if not wait:
            self._ck(self.lib.rca_peer_allgather(self.h, int(value), None))
            return None
        out = np.empty(self._peer_world, np.int64)
        self._ck(self.lib.rca_peer_allgather(self.h, int(value), out.ctypes.data))
        return out

    def peer_set_timeout(self, milliseconds: int):
        self._ck(self.lib.rca_peer_set_timeout(self.h, int(milliseconds)))

    def peer_wait(self):
        out = np.empty(self._peer_world, np.int64)
        self._ck(self.lib.rca_peer_wait(self.h, out.ctypes.data))
        return out

    # -- row gather over peer memory (include/rca_b200.h: rca_gather_*)
    def gather_window(self, words: int) -> bytes:
        """On the destination rank: allocate the window the ranks push their rows into; -> its IPC handle."""
        buf = C.create_string_buffer(self.PEER_HANDLE_BYTES)
        self._ck(self.lib.rca_gather_window(self.h, int(words), buf))
        return buf.raw

    def gather_connect(self, dst_rank: int, handle, words: int) -> None:
        self._ck(self.lib.rca_gather_connect(self.h, int(dst_rank), handle, int(words)))

    def gather_rows(self, result, level: int, offset_words: int = 0):
        """Collective: rows of `level` of `result` (an ExpandResult of this engine's last submission) from every
        rank into the destination's window at word `offset_words`.  -> (device pointer or 0, total rows, per-rank rows, this rank's push ms)."""
        assert result._gen == self._gen, "result was overwritten by a later expand()"
        self._ck(self.lib.rca_result_select(self.h, result.index))
        p, n, ms = C.c_void_p(), C.c_int64(), C.c_double()
        per = np.zeros(self._peer_world, np.int64)
        self._ck(self.lib.rca_gather_rows(self.h, int(level), int(offset_words), C.byref(p), C.byref(n), per.ctypes.data, C.byref(ms)))
        return (p.value or 0), int(n.value), per, float(ms.value)

    def match_pairs(self, msg_nodes, needles_a, needles_b):
        """The message filter of an instance query for many rows at once (generate_query.py:104-129):
        flag[k] = needles_a[k] in message(msg_nodes[k]) or needles_b[k] in message(msg_nodes[k]).
        msg_nodes: node ids whose messages were given to text_load(); needles: str, or None for 'no such text'."""
        n = len(msg_nodes)
        if n == 0:
            return np.zeros(0, dtype=bool)
        nodes = np.asarray(msg_nodes, dtype=np.int32)
        idx = np.searchsorted(self._text_nodes, nodes)
        if (idx >= len(self._text_nodes)).any() or (self._text_nodes[np.minimum(idx, len(self._text_nodes) - 1)] != nodes).any():
            raise ValueError("match_pairs: a node has no message in the loaded arena")
        uniq = {}
        na, nb = np.empty(n, np.int32), np.empty(n, np.int32)
        for col, arr in ((needles_a, na), (needles_b, nb)):
            for k, t in enumerate(col):
                arr[k] = -1 if t is None else uniq.setdefault(t, len(uniq))
        enc = [t.encode("utf-8") for t in uniq]
        off = np.zeros(len(enc) + 1, dtype=np.int64)
        np.cumsum([len(b) for b in enc], out=off[1:])
        blob = np.ascontiguousarray(np.frombuffer(b"".join(enc), dtype=np.uint8)) if off[-1] else np.zeros(1, np.uint8)
        idx = np.ascontiguousarray(idx, dtype=np.int32)
        out = np.empty(n, np.uint8)
        self._ck(self.lib.rca_text_match_pairs(self.h, blob.ctypes.data, off.ctypes.data, len(enc), idx.ctypes.data,
                                               na.ctypes.data, nb.ctypes.data, n, out.ctypes.data))
        return out.astype(bool)

    def state_lookup(self, nodes, ts, with_ids=True, ts_hi=None, state_kind=None):
        """HasState predicate for (entity node, timestamp ms) pairs -> (first, count, ids).
        With ts_hi: intervals with tmin <= ts and tmax > ts_hi (the reference's 'loose' form).
        state_kind: only STATE nodes of this kind (the statement's `(n2:KIND)` label)."""
        nodes = np.ascontiguousarray(nodes, dtype=np.int32)
        ts = np.ascontiguousarray(ts, dtype=np.int64)
        n = len(nodes)
        first = np.empty(n, np.int32)
        cnt = np.empty(n, np.int32)
        ids = np.empty((n, STATE_LIMIT), np.int32) if with_ids else None
        if state_kind is not None:
            hi = np.ascontiguousarray(ts if ts_hi is None else ts_hi, dtype=np.int64)
            self._ck(self.lib.rca_state_window_kind(self.h, nodes.ctypes.data, ts.ctypes.data, hi.ctypes.data, n, 0,
                                                    int(state_kind), first.ctypes.data, cnt.ctypes.data,
                                                    ids.ctypes.data if with_ids else None))
        elif ts_hi is None:
            self._ck(self.lib.rca_state_lookup(self.h, nodes.ctypes.data, ts.ctypes.data, n, 0, first.ctypes.data,
                                               cnt.ctypes.data, ids.ctypes.data if with_ids else None))
        else:
            hi = np.ascontiguousarray(ts_hi, dtype=np.int64)
            self._ck(self.lib.rca_state_window(self.h, nodes.ctypes.data, ts.ctypes.data, hi.ctypes.data, n, 0,
                                               first.ctypes.data, cnt.ctypes.data, ids.ctypes.data if with_ids else None))
        return first, cnt, ids

"""ctypes loader for oracle/csr_oracle.c - TEST INFRASTRUCTURE, NOT PRODUCT CODE.

PARITY UNPINNED (see cypher_oracle.py).  Takes a Snapshot's numpy arrays by pointer; results come
back as numpy arrays of rows, one array per path length."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libcsr_oracle.so")
MAX_HOPS = 8


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "csr_oracle.c")
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B", "libcsr_oracle.so"])
    return _LIB


def _host_tag() -> str:
    """Names the CPU this process runs on: a library built with -march=native must not travel to another host."""
    import hashlib
    try:
        with open("/proc/cpuinfo") as f:
            lines = [ln for ln in f if ln.startswith(("model name", "flags"))][:2]
    except OSError:
        lines = []
    return hashlib.sha1("".join(lines).encode()).hexdigest()[:12]


def build_native(force: bool = False) -> str:
    """The oracle built -O3 -march=native for THIS host (the timed CPU baseline of bench.py), in
    oracle/_native/<host tag>/; falls back to the portable build when the compiler refuses."""
    src = os.path.join(_HERE, "csr_oracle.c")
    rel = os.path.join("_native", _host_tag())
    lib = os.path.join(_HERE, rel, "libcsr_oracle.so")
    if force or not os.path.exists(lib) or os.path.getmtime(lib) < os.path.getmtime(src):
        try:
            subprocess.check_call(["make", "-s", "-C", _HERE, "-B", "native", f"NATIVE_DIR={rel}"])
        except (subprocess.CalledProcessError, OSError):
            return build(force)
    return lib


class _Graph(C.Structure):
    _fields_ = [("n_nodes", C.c_int32), ("n_kinds", C.c_int32),
                ("kind_start", C.c_void_p), ("out_rng", C.c_void_p), ("out_dst", C.c_void_p),
                ("in_rng", C.c_void_p), ("in_adj", C.c_void_p), ("rel_et", C.c_void_p),
                ("st_off", C.c_void_p), ("st_tmin", C.c_void_p), ("st_tmax", C.c_void_p),
                ("st_node", C.c_void_p)]


class _Varlen(C.Structure):
    _fields_ = [("direction", C.c_int32), ("min_hops", C.c_int32), ("max_hops", C.c_int32),
                ("entity_only", C.c_int32), ("node_simple", C.c_int32), ("dest_kind", C.c_int32),
                ("has_inter", C.c_int32), ("excl", C.c_void_p), ("inter", C.c_void_p),
                ("pos_kind", C.c_void_p)]


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class CsrOracle:
    def __init__(self, snap, lib_path=None):
        if lib_path is None:
            lib_path = build()
        self.lib = C.CDLL(lib_path)
        self.lib.orc_varlen.restype = C.c_int64
        self.lib.orc_chain.restype = C.c_int64
        self.lib.orc_varlen_step.restype = C.c_int64
        self.lib.orc_chain_step.restype = C.c_int64
        self.snap = snap
        names = ("kind_start", "out_rng", "out_dst", "in_rng", "in_adj", "rel_et", "st_off", "st_tmin",
                 "st_tmax", "st_node")
        self._keep = {n: np.ascontiguousarray(getattr(snap, n)) for n in names}
        self.g = _Graph(snap.n_nodes, snap.n_kinds, *[_ptr(self._keep[n]) for n in names])

    def max_threads(self) -> int:
        return int(self.lib.orc_max_threads())

    def _take(self, p, n_rows, width):
        if not p or n_rows == 0:
            return np.zeros((0, width), np.int32)
        a = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_int32)), shape=(n_rows * width,)).copy()
        self.lib.orc_free(C.c_void_p(p))
        return a.reshape(n_rows, width)

    def varlen(self, seeds, direction, max_hops, min_hops=1, dest_kind=-1, excl_kinds=(), inter_kinds=(),
               entity_only=False, node_simple=True, pos_kind=None, threads=0):
        """Returns ({length: rows[n, 2*length+1]}, scanned)."""
        nk = self.snap.n_kinds
        excl = np.zeros(nk, np.uint8)
        excl[list(excl_kinds)] = 1
        inter = np.zeros(nk, np.uint8)
        inter[list(inter_kinds)] = 1
        pk = None
        if pos_kind is not None:
            pk = np.ascontiguousarray(np.asarray(pos_kind, np.int32))
            assert len(pk) == max_hops + 1
        q = _Varlen({"out": 0, "in": 1, "und": 2}[direction], min_hops, max_hops, int(entity_only),
                    int(node_simple), dest_kind, int(len(inter_kinds) > 0), _ptr(excl), _ptr(inter),
                    _ptr(pk) if pk is not None else None)
        seeds = np.ascontiguousarray(np.asarray(seeds, np.int32))
        rows = (C.c_int64 * (MAX_HOPS + 1))()
        outs = (C.c_void_p * (MAX_HOPS + 1))()
        sc = C.c_int64(0)
        tot = self.lib.orc_varlen(C.byref(self.g), C.byref(q), _ptr(seeds), C.c_int64(len(seeds)), threads,
                                  rows, outs, C.byref(sc))
        if tot < 0:
            raise ValueError("bad arguments")
        res = {s: self._take(outs[s], rows[s], 2 * s + 1) for s in range(max_hops + 1) if rows[s]}
        return res, int(sc.value)

    def _varlen_desc(self, direction, max_hops, min_hops, dest_kind, excl_kinds, inter_kinds, entity_only, node_simple,
                     pos_kind):
        nk = self.snap.n_kinds
        excl = np.zeros(nk, np.uint8)
        excl[list(excl_kinds)] = 1
        inter = np.zeros(nk, np.uint8)
        inter[list(inter_kinds)] = 1
        pk = None
        if pos_kind is not None:
            pk = np.ascontiguousarray(np.asarray(pos_kind, np.int32))
            assert len(pk) == max_hops + 1
        q = _Varlen({"out": 0, "in": 1, "und": 2}[direction], min_hops, max_hops, int(entity_only),
                    int(node_simple), dest_kind, int(len(inter_kinds) > 0), _ptr(excl), _ptr(inter),
                    _ptr(pk) if pk is not None else None)
        return q, (excl, inter, pk)

    def reach(self, direction, max_hops, **kw):
        """Kind-level reachability table for the pruned step (orc_reach), or None when pruning does not apply."""
        q, keep = self._varlen_desc(direction, max_hops, kw.get("min_hops", 1), kw.get("dest_kind", -1),
                                    kw.get("excl_kinds", ()), kw.get("inter_kinds", ()), kw.get("entity_only", False),
                                    kw.get("node_simple", True), kw.get("pos_kind"))
        out = np.zeros(self.snap.n_kinds, np.uint32)
        return out if self.lib.orc_reach(C.byref(self.g), C.byref(q), _ptr(out)) else None

    def varlen_step(self, seeds, seed_ts, direction, max_hops, min_hops=1, dest_kind=-1, excl_kinds=(), inter_kinds=(),
                    entity_only=False, node_simple=True, pos_kind=None, reach=None, threads=0, keep_rows=True):
        """orc_varlen_step: the enumeration with the state columns computed in C (seed_ts given) and optional
        schema pruning (reach = self.reach(...)).  Returns ({length: rows}, scanned, total)."""
        q, keep = self._varlen_desc(direction, max_hops, min_hops, dest_kind, excl_kinds, inter_kinds, entity_only,
                                    node_simple, pos_kind)
        seeds = np.ascontiguousarray(np.asarray(seeds, np.int32))
        ts = None if seed_ts is None else np.ascontiguousarray(np.asarray(seed_ts, np.int64))
        rows = (C.c_int64 * (MAX_HOPS + 1))()
        outs = (C.c_void_p * (MAX_HOPS + 1))()
        sc = C.c_int64(0)
        tot = self.lib.orc_varlen_step(C.byref(self.g), C.byref(q), _ptr(seeds), _ptr(ts) if ts is not None else None,
                                       C.c_int64(len(seeds)), _ptr(reach) if reach is not None else None, threads,
                                       rows, outs, C.byref(sc))
        if tot < 0:
            raise ValueError("bad arguments")
        res = {}
        for s in range(max_hops + 1):
            if not rows[s]:
                continue
            w = 3 * s + 3 if ts is not None else 2 * s + 1
            if keep_rows:
                res[s] = self._take(outs[s], rows[s], w)
            else:
                self.lib.orc_free(C.c_void_p(outs[s]))
                res[s] = rows[s]
        return res, int(sc.value), int(tot)

    def chain_step(self, seeds, seed_ts, etypes, dirs, threads=0, keep_rows=True):
        """orc_chain_step: typed chain with the state columns computed in C.  Returns (rows or count, scanned)."""
        L = len(etypes)
        et = np.ascontiguousarray(np.asarray(etypes, np.int32))
        dr = np.ascontiguousarray(np.asarray(dirs, np.int32))
        seeds = np.ascontiguousarray(np.asarray(seeds, np.int32))
        ts = None if seed_ts is None else np.ascontiguousarray(np.asarray(seed_ts, np.int64))
        out = C.c_void_p()
        sc = C.c_int64(0)
        n = self.lib.orc_chain_step(C.byref(self.g), _ptr(et), _ptr(dr), L, _ptr(seeds),
                                    _ptr(ts) if ts is not None else None, C.c_int64(len(seeds)), threads,
                                    C.byref(out), C.byref(sc))
        if n < 0:
            raise ValueError("bad arguments")
        if keep_rows:
            return self._take(out.value, n, 3 * L + 3 if ts is not None else 2 * L + 1), int(sc.value)
        if out.value:
            self.lib.orc_free(out)
        return int(n), int(sc.value)

    def chain(self, seeds, etypes, dirs, threads=0):
        """Returns (rows[n, 2L+1], scanned)."""
        L = len(etypes)
        et = np.ascontiguousarray(np.asarray(etypes, np.int32))
        dr = np.ascontiguousarray(np.asarray(dirs, np.int32))
        seeds = np.ascontiguousarray(np.asarray(seeds, np.int32))
        out = C.c_void_p()
        sc = C.c_int64(0)
        n = self.lib.orc_chain(C.byref(self.g), _ptr(et), _ptr(dr), L, _ptr(seeds), C.c_int64(len(seeds)),
                               threads, C.byref(out), C.byref(sc))
        if n < 0:
            raise ValueError("bad arguments")
        return self._take(out.value, n, 2 * L + 1), int(sc.value)

    def states(self, nodes, ts, limit=10, ts_hi=None, threads=0):
        nodes = np.ascontiguousarray(np.asarray(nodes, np.int32))
        ts = np.ascontiguousarray(np.asarray(ts, np.int64))
        hi = ts if ts_hi is None else np.ascontiguousarray(np.asarray(ts_hi, np.int64))
        n = len(nodes)
        first = np.empty(n, np.int32)
        cnt = np.empty(n, np.int32)
        ids = np.empty((n, limit), np.int32)
        self.lib.orc_states(C.byref(self.g), _ptr(nodes), _ptr(ts), _ptr(hi), C.c_int64(n), limit, _ptr(first), _ptr(cnt),
                            _ptr(ids), int(threads))
        return first, cnt, ids

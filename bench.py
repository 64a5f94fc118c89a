#!/usr/bin/env python
"""bench.py - path-instances/s of the hot path on BASELINE.json's headline configuration.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

One step = one pass of the hot path over one incident batch (workloads.Config3): the undirected
<=4-hop Pod->Node enumeration from 100k seed Pods plus the five per-incident instance queries,
with the state predicate fused into emission.  Multi-GPU: graph replicated, one incident batch
per rank (weak scaling), all-gather of the per-rank instance counts; no data-path collective.

Printed line (rank 0): see the contract in the project brief.  `value` = instances of all ranks /
max-over-ranks device time with inputs resident in HBM; `e2e` = the same through host buffers
(pinned H2D of seeds and timestamps, D2H of every result row inside the timed region).

Beside the headline the line carries
  parity_check  untimed, after the timed legs and before the line is printed: ONE seed list (rank 0's batch) cut into contiguous
                work-balanced shards over the ranks, counts exchanged over peer memory AND over NCCL,
                rows gathered to rank 0 through the C ABI (rca_gather_rows) AND through NCCL, the
                gathered rows compared with rank 0's own expansion of the whole list and a seed
                sample with the CPU oracle.  A mismatch ends the run with a non-zero exit code.
  sharded       BASELINE config 4's shape as a strong-scaling measurement: one fixed job (10M-object
                graph, a million sorted seed Pods) cut over the ranks, with and without the payload
                gather to rank 0.
  config2       BASELINE config 2: 10k objects, every source kind -> Node, <= 4 hops, metapath ladder and instance
                expansion with kernel time and HBM fraction (rank 0).
  latency       BASELINE config 5: per-incident RCA graph latency.
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "path-instances/s"
FALLBACK_HBM_GBS = 6650.0


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


def csrc_digest():
    """Digest of the kernel sources: profiles/traffic.json names the digest it was captured at, and a capture
    that no longer matches the sources is not quoted."""
    h = hashlib.sha1()
    d = os.path.join(ROOT, "k8s-llm-rca_b200", "csrc")
    for f in sorted(os.listdir(d)) + [os.path.join(ROOT, "include", "rca_b200.h")]:
        if f.endswith((".cu", ".cuh", ".h")):
            with open(f if os.path.isabs(f) else os.path.join(d, f), "rb") as fh:
                h.update(fh.read())
    return h.hexdigest()[:16]


def captured_traffic(key):
    """(bytes per launch set or None, note) from profiles/traffic.json, only if it was captured from these sources."""
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(tp):
        return None, "no ncu capture committed"
    try:
        d = json.load(open(tp))
    except Exception:
        return None, "profiles/traffic.json unreadable"
    if d.get("csrc_sha") != csrc_digest():
        return None, f"capture is from other kernel sources (captured at {d.get('csrc_sha')}, commit {d.get('commit')})"
    return d.get(key), f"ncu --set full capture at commit {d.get('commit')} ({d.get('source')})"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 50 ms.  It is started before the warm-up
    (nvidia-smi needs a few hundred ms to produce its first line) and the samples are cut to the
    timed window by their timestamps."""
    Q = ("timestamp,clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.t0, self.t1 = index, [], None, None, None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def window(self, t0, t1):
        self.t0, self.t1 = t0, t1

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)                      # let the last sample of the window arrive
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.t.join(timeout=2)
        rows = [r for (t, r) in self.rows if len(r) >= 7 and (self.t0 is None or self.t0 - 0.05 <= t <= self.t1 + 0.15)]
        in_window = True
        if not rows:
            rows, in_window = [r for (_, r) in self.rows if len(r) >= 7], False
        sm = [float(r[1]) for r in rows if r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in rows if r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in rows for i in range(4) if r[3 + i].lower() == "active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm), "samples_in_timed_window": in_window}


def cpu_arm(w, pods, ev, ts, steps, warm):
    """The CPU restatement of the reference's path (oracle/csr_oracle.c, kind "port" - Neo4j itself is unavailable
    on this box) on the WHOLE batch the GPU arm runs: every step is one set of C calls (enumeration, the five
    instance queries, the state predicate per node of every row), built -O3 -march=native for this host, all host
    threads.  Two figures: the restatement as the reference's statements read (no pruning - Neo4j's variable-length
    expand evaluates the path predicates after expanding), and the same job with the kind-level schema pruning the
    engine's plans use - the stronger CPU program, reported beside it."""
    from oracle import csr
    orc = csr.CsrOracle(w.snap, csr.build_native())
    threads = os.cpu_count() or orc.max_threads()        # torchrun exports OMP_NUM_THREADS=1: ask for the cores explicitly
    n = len(pods)
    k = max(1, n // 20)
    out = {}
    for name, pruned in (("plain", False), ("pruned", True)):
        for _ in range(max(1, warm)):
            w.oracle_step(orc, pods[:k], ev[:k], ts[:k], threads=threads, pruned=pruned)
        t0 = time.perf_counter()
        inst = 0
        for _ in range(steps):
            tot, _ = w.oracle_step(orc, pods, ev, ts, threads=threads, pruned=pruned)
            inst += tot
        dt = time.perf_counter() - t0
        out[name] = (inst / dt, dt / steps, inst // steps)
    return out, threads


def run_reference(args, rank, world):
    """--impl reference: see cpu_arm.  Rank 0 alone runs it."""
    if rank != 0:
        return
    from k8s_llm_rca_b200 import workloads
    w = workloads.Config3(n_objects=args.objects, n_events=args.events)
    pods, ev, ts = w.batch(0)
    res, threads = cpu_arm(w, pods, ev, ts, args.steps, min(args.warmup, 2))
    v, sec, inst = res["plain"]
    sample = f"all {len(pods)} incidents of the config-3 batch per step ({inst} instances, {sec:.2f} s per step)"
    emit({
        "metric": METRIC, "value": v, "unit": "instances/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * sec, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int32", "data": "synthetic", "impl": "reference",
        "config": {"workload": w.name, "objects": args.objects, "incidents_per_step": int(len(pods)),
                   "instances_per_step": int(inst), "max_hops": w.max_hops},
        "cpu_baseline": {"value": v, "unit": "instances/s", "cores": threads, "kind": "port", "sample": sample,
                         "pruned": {"value": res["pruned"][0], "ms_per_step": 1e3 * res["pruned"][1],
                                    "what": "the same C restatement with kind-level schema pruning (result-neutral)"},
                         "build": "gcc -O3 -march=native -fopenmp, state predicate in C, one call set per step"},
        "e2e": {"value": v, "unit": "instances/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


_REAL_STDOUT = None


def emit(obj):
    """The one JSON line of this run, on the process's real stdout."""
    data = (json.dumps(obj) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def sort_rows_t(t):
    """Rows of an int32 tensor in lexicographic order (stable sorts, last column first) - on the device."""
    import torch
    for c in range(t.shape[1] - 1, -1, -1):
        t = t[torch.sort(t[:, c], stable=True).indices]
    return t


def parity_leg(w, eng, rank, world, dev, peer_ok, oracle_sample):
    """Untimed: the seed-PARTITIONED path checked end to end.  Returns the dict for the JSON line (rank 0)."""
    import torch
    import torch.distributed as dist
    from k8s_llm_rca_b200 import parallel
    pods, ev, ts = w.batch(0)                       # ONE seed list for all ranks: rank 0's batch
    order = np.argsort(pods, kind="stable")
    pods, ts = pods[order], ts[order]               # sorted: rank-order concatenation is then ordered by seed
    L = w.max_hops
    my_pods, my_ts, (lo, hi) = parallel.shard_seeds(w.snap, pods, rank, world, "und", True, seed_ts=ts)
    res = eng.expand(w.varlen, my_pods, my_ts)
    info = {"ranks": world, "seeds": int(len(pods)), "shard": [int(lo), int(hi)], "ok": False}
    fails = []
    # counts: peer memory vs NCCL
    nccl = parallel.allgather_counts(res.emitted, device=dev)             # [world, L+1]
    if world > 1 and peer_ok:
        for s in range(1, L + 1):
            got = eng.peer_allgather(res.emitted[s])
            if got.tolist() != nccl[:, s].tolist():
                fails.append(f"level {s}: peer count exchange {got.tolist()} != NCCL {nccl[:, s].tolist()}")
        info["counts_peer_eq_nccl"] = not fails
    totals = nccl.sum(dim=0).tolist()
    info["rows"] = int(sum(totals))
    # rows: C-ABI gather over peer memory vs NCCL gather, every level, into one window
    gathered = {}
    if world > 1 and peer_ok:
        offs, at = {}, 0
        for s in range(1, L + 1):
            offs[s] = at
            at += totals[s] * res.row_len[s]
        parallel.connect_gather(eng, at + 16, dst=0)
        for s in range(1, L + 1):
            if not totals[s]:
                continue
            got, per, _ = parallel.gathered_rows(eng, res, s, dst=0, offset_words=offs[s])
            ref = parallel.gather_rows(parallel.device_rows(res, s), nccl[:, s], dst=0)
            if rank == 0:
                if not torch.equal(got, ref):
                    fails.append(f"level {s}: rca_gather_rows differs from the NCCL gather")
                gathered[s] = got.clone()
        info["gather_peer_eq_nccl"] = not fails
    else:
        for s in range(1, L + 1):
            if totals[s]:
                ref = parallel.gather_rows(parallel.device_rows(res, s), nccl[:, s], dst=0)
                if rank == 0:
                    gathered[s] = ref.clone()
    if rank == 0:
        # the gathered rows against this rank's own expansion of the WHOLE list (same multiset, row for row)
        d_p, d_t = torch.from_numpy(pods).to(dev), torch.from_numpy(ts).to(dev)
        torch.cuda.synchronize()
        whole = eng.expand(w.varlen, d_p, d_t)
        for s in range(1, L + 1):
            if whole.emitted[s] != totals[s]:
                fails.append(f"level {s}: {totals[s]} rows over the shards, {whole.emitted[s]} from one GPU")
                continue
            if totals[s] and not torch.equal(sort_rows_t(gathered[s]), sort_rows_t(parallel.device_rows(whole, s).clone())):
                fails.append(f"level {s}: sharded rows differ from the single-GPU rows")
        info["sharded_eq_single_gpu"] = not any("single" in f or "one GPU" in f for f in fails)
        # a seed sample against the CPU oracle, state columns included
        from oracle import csr
        orc = csr.CsrOracle(w.snap)
        uniq = np.unique(pods)                       # a Pod may have more than one incident: take all of a chosen Pod's
        vals = uniq[np.unique(np.linspace(0, len(uniq) - 1, min(oracle_sample, len(uniq))).astype(np.int64))]
        mask = np.isin(pods, vals)
        sp, st = pods[mask], ts[mask]
        want, _, _ = orc.varlen_step(sp, st, "und", L, dest_kind=w.snap.kind_index("Node"), excl_kinds=w.excl,
                                     entity_only=True)
        d_sp = torch.from_numpy(sp).to(dev)
        n_cmp = 0
        for s in range(1, L + 1):
            if not totals[s]:
                if s in want:
                    fails.append(f"level {s}: oracle has rows, the GPUs none")
                continue
            g = gathered[s]
            g = g[torch.isin(g[:, 0], d_sp)]
            a = sort_rows_t(g).cpu().numpy()
            b = want.get(s, np.zeros((0, res.row_len[s]), np.int32))
            b = b[np.lexsort(b.T[::-1])] if len(b) else b
            n_cmp += len(b)
            if a.shape != b.shape or not np.array_equal(a, b):
                fails.append(f"level {s}: rows of the seed sample differ from the CPU oracle")
        info["oracle_sample"] = {"seeds": int(len(sp)), "rows": int(n_cmp)}
        info["ok"] = not fails
        if fails:
            info["failures"] = fails[:6]
    flag = torch.tensor([0 if (rank == 0 and fails) else 1], device=dev)
    if world > 1:
        dist.broadcast(flag, src=0)
    info["ok"] = bool(flag.item()) and not fails
    return info


def config2_leg(local_rank, peak):
    """BASELINE config 2: a 10k-object cluster, every source kind -> Node, <= 4 hops, the full ladder of
    find_metapath (directed, undirected, one hop, via Namespace; shortest length kept) - on the metagraph, as the
    reference runs it, through run_query; and the same statement shape over the INSTANCE graph (all nodes of the kind
    as seeds, directed then undirected), which is where the kernels have work.  Small and latency-bound by nature:
    reported for completeness, with its kernel time and the HBM fraction its algorithmic bytes amount to."""
    import contextlib
    import io
    from k8s_llm_rca_b200 import dropin, engine as E, executor, synth
    snap = synth.generate(10_000, 1000, synth.config_seed(1))
    executor.register_snapshot("config2", snap)
    mx = executor.GpuQueryExecutor("snapshot://config2#metagraph", device=local_rank)
    dropin.install()
    from find_metapath.find_srckind_metapath_neo4j import find_metapath
    eng = E.Engine(snap, device=local_rank)
    K = snap.kind_index
    node, excl = K("Node"), [K("Event"), K("Namespace")]
    out = {"objects": 10_000, "max_hops": 4, "dest": "Node", "metapaths": {}, "instances": {}}
    kernel_ms = bytes_alg = total = 0
    t0 = time.perf_counter()
    for k, kind in enumerate(snap.kinds):
        if kind.is_state or kind.name in ("Event", "Namespace", "Node"):
            continue
        with contextlib.redirect_stdout(io.StringIO()):
            try:
                mps = find_metapath(mx, kind.name, "Node", [])
            except ValueError:                                     # no rung of the ladder finds a path (reference :153)
                mps = []
        out["metapaths"][kind.name] = {"count": len(mps), "length": len(mps[0]) if mps else None}
        seeds = snap.nodes_of_kind(k)
        n_kind = 0
        for direction in ("out", "und"):
            r = eng.expand(E.varlen_plan(direction, 4, dest_kind=node, excl_kinds=excl, entity_only=True, pos_kind=[k]), seeds)
            kernel_ms += r.kernel_ms
            bytes_alg += r.algorithmic_bytes()
            n_kind += r.total
        out["instances"][kind.name] = n_kind
        total += n_kind
    out.update({"wall_ms": (time.perf_counter() - t0) * 1e3, "kernel_ms": kernel_ms, "instances_total": total,
                "algorithmic_GBps": bytes_alg / (kernel_ms * 1e-3) / 1e9 if kernel_ms else None,
                "hbm_frac": bytes_alg / (kernel_ms * 1e-3) / 1e9 / peak if kernel_ms else None,
                "note": "launch- and latency-bound: a few thousand seeds per launch; the metapath sets equal the oracle's (tests)"})
    eng.close()
    mx.close()
    return out


def sharded_leg(args, rank, world, local_rank, dev, barrier):
    """BASELINE config 4's shape, strong scaling: one fixed job over `world` GPUs."""
    import torch
    import torch.distributed as dist
    from k8s_llm_rca_b200 import engine as E
    from k8s_llm_rca_b200 import parallel, workloads
    t0 = time.time()
    w4 = workloads.Config4(n_objects=args.sharded_objects, n_seeds=args.sharded_seeds, rank=rank, barrier=barrier)
    t_graph = time.time() - t0
    eng = E.Engine(w4.snap, device=local_rank)
    stream = torch.cuda.Stream(priority=-1)
    eng.set_stream(stream.cuda_stream)
    seeds, _, (lo, hi) = parallel.shard_seeds(w4.snap, w4.seeds, rank, world, "und", True)
    steps = max(2, min(args.steps, args.sharded_steps))
    out = {"workload": w4.name, "objects": args.sharded_objects, "seeds": int(len(w4.seeds)), "steps": steps,
           "graph_bytes_in_hbm": int(eng.graph_bytes()), "graph_ready_s": round(t_graph, 1)}
    with torch.cuda.stream(stream):
        d_seeds = torch.from_numpy(np.ascontiguousarray(seeds)).to(dev)
        torch.cuda.synchronize()
        peer_ok = False
        if world > 1:
            try:
                parallel.connect_peers(eng)
                eng.peer_set_timeout(20_000)
                peer_ok = eng.peer_allgather(rank + 1).tolist() == list(range(1, world + 1))
            except Exception as ex:
                print(f"bench.py: sharded leg: peer memory unavailable ({ex})", file=sys.stderr)
            flag = torch.tensor([1 if peer_ok else 0], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            peer_ok = bool(flag.item())

        def timed(fn):
            for _ in range(2):
                fn()
            barrier()
            ms_kernel, e0, e1 = 0.0, torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            last = None
            for _ in range(steps):
                last = fn()
                ms_kernel += last[1]
            e1.record(stream)
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / steps, ms_kernel / steps, last[0]

        def reduce_ms(ms, kernel_ms, inst):
            t = torch.tensor([ms, kernel_ms, -kernel_ms], dtype=torch.float64, device=dev)
            n = torch.tensor([inst], dtype=torch.int64, device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                dist.all_reduce(n, op=dist.ReduceOp.SUM)
            return float(t[0]), float(t[1]), float(-t[2]), int(n[0])

        # (1) <= 5 hops, counted; the count exchange is the only thing that crosses NVLink
        def count5():
            r = eng.expand(w4.count5, d_seeds)
            if world > 1 and peer_ok:
                eng.peer_allgather(r.total)
            return r.total, r.kernel_ms
        ms, kmax, kmin, inst = reduce_ms(*timed(count5))
        out["count5"] = {"instances": inst, "ms_per_step": ms, "value": inst / (ms * 1e-3), "unit": "instances/s",
                         "rank_kernel_ms": {"max": kmax, "min": kmin}}

        # (2) <= 4 hops with rows: (a) rows stay on their ranks, (b) gathered to rank 0's window
        def rows4():
            r = eng.expand(w4.rows4, d_seeds)
            if world > 1 and peer_ok:
                eng.peer_allgather(r.total)
            return r.total, r.kernel_ms, r
        r0 = rows4()[2]
        cnt = parallel.allgather_counts(r0.emitted, device=dev)
        totals = cnt.sum(dim=0).tolist()
        ms, kmax, kmin, inst = reduce_ms(*timed(lambda: rows4()[:2]))
        out["rows4"] = {"instances": inst, "ms_per_step": ms, "value": inst / (ms * 1e-3), "unit": "instances/s",
                        "rank_kernel_ms": {"max": kmax, "min": kmin}}
        if world > 1 and peer_ok:
            offs, at = {}, 0
            for s in range(1, 5):
                offs[s] = at
                at += totals[s] * r0.row_len[s]
            parallel.connect_gather(eng, at + 16, dst=0)
            push = [0.0]

            def rows4_gather():
                tot, kms, r = rows4()
                for s in range(1, 5):
                    if totals[s]:
                        _, _, _, pms = eng.gather_rows(r, s, offs[s])
                        push[0] += pms
                return tot, kms
            push[0] = 0.0
            ms_g, _, _, inst_g = reduce_ms(*timed(rows4_gather))
            words_in = sum((totals[s] - int(cnt[0, s])) * r0.row_len[s] for s in range(1, 5))     # what crosses NVLink into rank 0
            gather_ms = max(ms_g - ms, 1e-6)
            out["rows4_gathered"] = {"instances": inst_g, "ms_per_step": ms_g, "value": inst_g / (ms_g * 1e-3),
                                     "unit": "instances/s", "bytes_into_rank0_per_step": int(4 * words_in),
                                     "gather_ms_per_step": gather_ms,
                                     "gather_GBps_into_rank0": 4 * words_in / (gather_ms * 1e-3) / 1e9,
                                     "how": "rca_gather_rows: every rank stores its rows into rank 0's window over NVLink "
                                            "peer memory (16-byte remote stores), offsets from the count exchange"}
            # the gathered level 4 is what one GPU produces for the whole list (counts; rows are checked by parity_check)
        out["limiter_note"] = ("strong scaling: per-rank kernel time falls with the shard; what does not shrink is the fixed "
                               "cost of a submission (uploads, launches, one synchronisation) and, with the payload gather, "
                               "rank 0's NVLink ingress")
    eng.close()
    return out


def main():
    # libraries (NCCL's version banner, for one) write to file descriptor 1; the contract is ONE JSON line there,
    # so everything else that lands on fd 1 goes to stderr and the line is written to the saved descriptor
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--objects", type=int, default=1_000_000)
    ap.add_argument("--events", type=int, default=108_696)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--latency-incidents", type=int, default=10_000,
                    help="config 5: incidents replayed as one batch (drawn with replacement); 0 skips the latency leg")
    ap.add_argument("--latency-singles", type=int, default=200, help="incidents replayed one by one for p50/p99")
    ap.add_argument("--parity-sample", type=int, default=2000, help="seeds of the parity leg compared with the CPU oracle; 0 skips the leg")
    ap.add_argument("--sharded-objects", type=int, default=10_000_000, help="config 4 shape, strong scaling; 0 skips the leg")
    ap.add_argument("--sharded-seeds", type=int, default=1_000_000)
    ap.add_argument("--sharded-steps", type=int, default=10)
    ap.add_argument("--config2", type=int, default=1, help="0 skips the config-2 leg (10k objects, every source kind -> Node)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the b200 arm has no CPU path (use --impl reference)")
    torch.cuda.set_device(local_rank)
    if world > 1 and os.environ.get("RCA_BENCH_PIN", "1") == "1":
        # one process per GPU shares the host with the others: keep each rank on its own pair of cores, so that the
        # host side of a step (about 0.15 ms of launches per 0.5 ms step) is not migrated or preempted by a neighbour
        try:
            cores = sorted(os.sched_getaffinity(0))
            per = len(cores) // world
            if per >= 1:
                os.sched_setaffinity(0, set(cores[local_rank * per:(local_rank + 1) * per]))
        except (AttributeError, OSError):
            pass
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    from k8s_llm_rca_b200 import engine as E
    from k8s_llm_rca_b200 import workloads
    w = workloads.Config3(n_objects=args.objects, n_events=args.events)
    eng = E.Engine(w.snap, device=local_rank)
    stream = torch.cuda.Stream(priority=-1)      # above the engine's side streams (include/rca_b200.h: rca_set_stream)
    eng.set_stream(stream.cuda_stream)
    pods, ev, ts = w.batch(rank)
    dev = torch.device("cuda", local_rank)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    parity = None
    with torch.cuda.stream(stream):
        d_pods, d_ev, d_ts = (torch.from_numpy(x).to(dev) for x in (pods, ev, ts))
        h_pods, h_ev, h_ts = (torch.from_numpy(x).pin_memory() for x in (pods, ev, ts))
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)       # > 126 MB L2
        counts = torch.zeros(world, dtype=torch.int64, device=dev)
        my_count = torch.zeros(1, dtype=torch.int64, device=dev)
        peer_ok = False
        last_counts = [None]
        if world > 1:
            from k8s_llm_rca_b200 import parallel
            try:
                parallel.connect_peers(eng)
                eng.peer_set_timeout(20_000)         # the parity leg gives rank 0 seconds of extra work
                got = eng.peer_allgather(rank + 1)
                peer_ok = got.tolist() == list(range(1, world + 1))
            except Exception as ex:                                              # no peer access between these GPUs: NCCL does it
                print(f"bench.py: peer-memory count exchange unavailable ({ex}); using NCCL all-gather", file=sys.stderr)
            flag = torch.tensor([1 if peer_ok else 0], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)                          # all ranks take the same road
            peer_ok = bool(flag.item())

        # size the pinned result buffer from one untimed run
        tot0, res0 = w.run(eng, d_pods, d_ev, d_ts)
        max_words = max(max(r.emitted[s] * r.row_len[s] for s in range(r.levels + 1)) for _, r in res0)
        h_rows = torch.empty(max_words + 16, dtype=torch.int32).pin_memory().numpy()
        d2h = [0]

        def fetch(r):
            for s in range(1, r.levels + 1):
                if r.emitted[s]:
                    r.rows(s, out=h_rows)
                    d2h[0] += r.emitted[s] * r.row_len[s] * 4

        def step(resident):
            if resident:
                tot, res = w.run(eng, d_pods, d_ev, d_ts)
            else:
                tot, res = w.run(eng, h_pods.numpy(), h_ev.numpy(), h_ts.numpy(), fetch=fetch)
            if world > 1:       # the only collective of the path: every rank learns every rank's count
                if peer_ok:     # ... over peer memory: one small kernel on the engine's stream, nothing on the host
                    eng.peer_allgather(tot, wait=False)
                else:
                    my_count.fill_(tot)
                    dist.all_gather_into_tensor(counts, my_count)
            return tot, res

        def timed(resident, steps):
            evs = []
            inst, launches, per_level = 0, 0, {}
            for it in range(steps):
                flush.zero_()                                                # evict the graph from L2
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                tot, res = step(resident)
                if peer_ok and it == steps - 1:                              # the exchanges run one step deferred: the last
                    last_counts[0] = eng.peer_wait()                         # one is collected inside the timed region
                e1.record(stream)
                evs.append((e0, e1))
                inst += tot
                for name, r in res:
                    launches += r.launches
                    for s in range(1, r.levels + 1):
                        a = per_level.setdefault((name, s), [0.0, 0, 0])
                        a[0] += r.level_ms[s]
                        a[1] += r.level_bytes(s)
                        a[2] += 1
            torch.cuda.synchronize()
            ms = sum(a.elapsed_time(b) for a, b in evs)
            return ms, inst, launches, per_level

        def timed_e2e_pipelined(steps, wl=None):
            """End to end with the copy-out of step k travelling while step k+1 runs: the engine alternates between
            two output arenas (rca_set_double_buffer) and the rows leave through rca_result_copy_async.  Every step
            still uploads its seeds from pinned host memory and downloads all of its rows; ONE pair of events
            brackets all steps (L2 flushes included).  wl: the workload (default: the headline one)."""
            wl = wl or w
            sized = res0 if wl is w else wl.run(eng, d_pods, d_ev, d_ts)[1]
            total_words = sum(r.emitted[s] * r.row_len[s] for _, r in sized for s in range(1, r.levels + 1))
            h_all = torch.empty(total_words + 64, dtype=torch.int32).pin_memory().numpy()
            cursor = [0]

            def fetch_async(r):
                if r.index == 0:
                    eng.result_wait()                    # the previous step's rows have arrived: the buffer is free again
                    cursor[0] = 0
                for s in range(1, r.levels + 1):
                    n = r.emitted[s] * r.row_len[s]
                    if n:
                        r.rows_async(s, h_all[cursor[0]:cursor[0] + n])
                        cursor[0] += n
                        d2h[0] += n * 4

            def one():
                tot, _ = wl.run(eng, h_pods.numpy(), h_ev.numpy(), h_ts.numpy(), fetch=fetch_async)
                if world > 1:
                    if peer_ok:
                        eng.peer_allgather(tot, wait=False)
                    else:
                        my_count.fill_(tot)
                        dist.all_gather_into_tensor(counts, my_count)
                return tot

            eng.set_double_buffer(True)
            for _ in range(max(args.warmup, 3)):         # both arenas reach their size before the clock starts
                one()
            eng.result_wait()
            barrier()
            d2h[0] = 0
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            e0.record(stream)
            inst = 0
            for it in range(steps):
                flush.zero_()
                inst += one()
            eng.result_wait()
            if peer_ok:
                last_counts[0] = eng.peer_wait()
            e1.record(stream)
            torch.cuda.synchronize()
            wall = time.perf_counter() - t0
            eng.set_double_buffer(False)
            return e0.elapsed_time(e1), inst, wall

        sampler = ClockSampler(local_rank)
        if rank == 0:
            sampler.start()
        for _ in range(args.warmup):
            step(True)
        for _ in range(args.warmup):
            step(False)
        barrier()
        t_win0 = time.time()
        ms_res, inst_res, launches, per_level = timed(True, args.steps)
        barrier()
        d2h[0] = 0
        ms_serial, inst_serial, _, _ = timed(False, args.steps)     # one step after the other: copy-out, then the next step
        barrier()
        ms_e2e, inst_e2e, wall_e2e = timed_e2e_pipelined(args.steps)
        d2h_full = d2h[0]
        barrier()
        # the same end to end with mask-only rows (RCA_STATES_MASK): the missing-state mask instead of a state id per node
        w_mask = workloads.Config3(snap=w.snap, with_states="mask")
        ms_e2e_m, inst_e2e_m, _ = timed_e2e_pipelined(args.steps, w_mask)
        d2h_mask, d2h[0] = d2h[0], d2h_full
        barrier()
        # what the platform gives a plain device->host copy when every rank copies at once: the ceiling of any e2e figure
        probe_words = min(max_words, 64 << 20)
        src = torch.empty(probe_words, dtype=torch.int32, device=dev)
        dst = torch.from_numpy(h_rows[:probe_words])
        for _ in range(2):
            dst.copy_(src, non_blocking=True)
        barrier()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record(stream)
        for _ in range(8):
            dst.copy_(src, non_blocking=True)
        c1.record(stream)
        torch.cuda.synchronize()
        d2h_ceiling = 8 * probe_words * 4 / (c0.elapsed_time(c1) * 1e-3) / 1e9
        del src
        if peer_ok:                                                              # what the last exchange delivered
            last = last_counts[0]
            assert len(last) == world and (last > 0).all(), f"count exchange returned {last}"
        barrier()
        sampler.window(t_win0, time.time())
        clocks = sampler.stop() if rank == 0 else None

    # max over ranks of the device time, sum over ranks of the instances
    red = torch.tensor([ms_res, ms_e2e, ms_serial, ms_e2e_m], dtype=torch.float64, device=dev)
    tot = torch.tensor([inst_res, inst_e2e, inst_e2e_m], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(red, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    ms_res, ms_e2e, ms_serial, ms_e2e_m = red.tolist()
    inst_res, inst_e2e, inst_e2e_m = tot.tolist()

    # ---- config 5: per-incident RCA graph latency (LLM answers recorded), one by one and batched
    latency = None
    if args.latency_incidents > 0:
        from k8s_llm_rca_b200 import executor, replay
        executor.register_snapshot("bench", w.snap)
        sx = executor.GpuQueryExecutor("snapshot://bench", device=local_rank)
        mx = executor.GpuQueryExecutor("snapshot://bench#metagraph", device=local_rank)
        rp = replay.IncidentReplayer(sx, mx)
        rng = np.random.default_rng(5)
        msgs = w.snap.props.msgs
        draw = [msgs[i] for i in rng.integers(0, len(msgs), size=args.latency_incidents)]
        mine = draw[rank::world]
        rp.run_batch_arrays(mine)                                 # warm-up: message arena upload, metapath and name-arena caches
        barrier()
        t0 = time.perf_counter()
        out = rp.run_batch_arrays(mine)
        torch.cuda.synchronize()
        bt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        done = out.done_ms[out.done_ms > 0]
        pct = torch.tensor([float(np.percentile(done, 50)) if len(done) else 0.0, float(np.percentile(done, 99)) if len(done) else 0.0],
                           dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(bt, op=dist.ReduceOp.MAX)
            dist.all_reduce(pct, op=dist.ReduceOp.MAX)
        if rank == 0:
            lat = []
            for m in draw[:args.latency_singles]:
                t1 = time.perf_counter()
                rp.run_one(m)
                lat.append((time.perf_counter() - t1) * 1e3)
            lat = np.array(lat)
            latency = {"single_p50_ms": float(np.percentile(lat, 50)), "single_p99_ms": float(np.percentile(lat, 99)),
                       "single_n": int(len(lat)), "batch_incidents": args.latency_incidents,
                       "batch_total_ms": float(bt) * 1e3, "batch_ms_per_incident": float(bt) * 1e3 / args.latency_incidents,
                       "batch_p50_ms": float(pct[0]), "batch_p99_ms": float(pct[1]),
                       "batch_percentiles_are": "time from the start of the batch to the completion of the incident's "
                                                "(source, destination) group, max over ranks",
                       "rows_in_batch": out.n_rows,
                       "what": "per incident: CONTAINS seed lookup + source kind + metapath ladder + instance query + "
                               "message filter + state check per node, LLM answers recorded; single = through "
                               "run_query one statement at a time, batch = one launch set for all incidents"}
        sx.close()
        mx.close()
    # ---- config 4's shape: strong scaling of one fixed job, with and without the payload gather
    sharded = None
    if args.sharded_objects > 0:
        barrier()
        sharded = sharded_leg(args, rank, world, local_rank, dev, barrier)
        barrier()

    # ---- parity of the seed-partitioned path: untimed, after the timed legs (so that nothing it sets up - NCCL
    # point-to-point channels, the gather window, rank 0's oracle - can colour the headline), before the line is printed
    if args.parity_sample > 0:
        with torch.cuda.stream(stream):
            torch.cuda.synchronize()
            parity = parity_leg(w, eng, rank, world, dev, peer_ok, args.parity_sample)
        barrier()
        if not parity["ok"]:
            if rank == 0:
                emit({"metric": METRIC, "value": None, "n_gpus": world, "parity_check": parity,
                      "error": "the seed-partitioned path does not reproduce the single-GPU / oracle rows"})
            raise SystemExit(3)
    graph_bytes = int(eng.graph_bytes())
    eng.close()
    del d_pods, d_ev, d_ts, flush
    torch.cuda.empty_cache()

    if rank == 0:
        peak, peak_src = peaks()
        config2 = config2_leg(local_rank, peak) if args.config2 else None
        # dominant kernel = the level launch with the largest share of the step
        (kname, ks), (kms, kbytes, kcalls) = max(per_level.items(), key=lambda kv: kv[1][0])
        achieved = (kbytes / kcalls) / (kms / kcalls * 1e-3) / 1e9
        traffic, traffic_note = captured_traffic(f"{kname}/level{ks}")
        dram_frac = (traffic / (kms / kcalls * 1e-3) / 1e9 / peak) if traffic else None
        out = {
            "metric": METRIC, "value": inst_res / (ms_res * 1e-3), "unit": "instances/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_res / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": {"workload": w.name, "objects": args.objects, "incidents_per_step_per_gpu": int(len(pods)),
                       "instances_per_step": inst_res // args.steps, "max_hops": w.max_hops,
                       "graph_bytes_in_hbm": graph_bytes,
                       "l2": f"flushed between timed steps (256 MiB write); graph ({graph_bytes / 1e6:.0f} MB) also exceeds the 126 MB L2",
                       "parallelism": f"graph replicated x{world}, one incident batch per GPU, all-gather of counts"
                                      + (" over peer memory (rca_peer_allgather)" if peer_ok else " (NCCL)" if world > 1 else "")},
            "e2e": {"value": inst_e2e / (ms_e2e * 1e-3), "unit": "instances/s",
                    "h2d_bytes_per_step": int(pods.nbytes + ev.nbytes + 2 * ts.nbytes),     # the five instance queries share one seed list
                    "d2h_bytes_per_step": int(d2h[0] // args.steps), "ms_per_step": ms_e2e / args.steps,
                    "wall_ms_per_step": 1e3 * wall_e2e / args.steps,
                    "d2h_GBps_per_rank": (d2h[0] / args.steps) / (ms_e2e / args.steps * 1e-3) / 1e9,
                    "d2h_ceiling_GBps_per_rank": d2h_ceiling,
                    "d2h_ceiling_is": f"plain pinned device->host copies of {probe_words * 4 >> 20} MiB, all {world} ranks copying at once (rank 0's figure)",
                    "how": "every step: seeds host->device from pinned memory, rca_expand_many, ALL rows device->host into "
                           "pinned memory; the copy-out of a step overlaps the next step (two output arenas, "
                           "rca_result_copy_async); one pair of CUDA events around all steps, L2 flushes included",
                    "serial_ms_per_step": ms_serial / args.steps,
                    "mask_rows": {"value": inst_e2e_m / (ms_e2e_m * 1e-3), "unit": "instances/s",
                                  "ms_per_step": ms_e2e_m / args.steps, "d2h_bytes_per_step": int(d2h_mask // args.steps),
                                  "d2h_GBps_per_rank": (d2h_mask / args.steps) / (ms_e2e_m / args.steps * 1e-3) / 1e9,
                                  "what": "the same step with with_states = RCA_STATES_MASK: rows carry the missing-state mask, "
                                          "not a state id per node (12 -> 8 words per three-hop row); same copies, same timing"}},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "dram_frac": dram_frac, "traffic_source": traffic_note,
                         "kernel": f"level {ks} of the {kname} plan (its count + scan + fill launches)",
                         "peak_source": peak_src, "bytes_per_launch": kbytes // kcalls,
                         "ms_per_launch": kms / kcalls, "share_of_step": (kms / kcalls) / (ms_res / args.steps),
                         "frac_is": "SURVEY 8(d)'s algorithmic bytes / time / peak; dram_frac = ncu DRAM bytes / time / peak"},
            "clocks": clocks,
            "parity_check": parity,
            "config2": config2,
            "sharded": sharded,
            "latency": latency,
        }
        if world == 1 and not args.no_cpu_baseline:
            res, cores = cpu_arm(w, pods, ev, ts, 1, 1)
            v, sec, inst = res["plain"]
            out["cpu_baseline"] = {"value": v, "unit": "instances/s", "cores": cores, "kind": "port",
                                   "sample": f"all {len(pods)} incidents of the same batch, one pass, {sec:.1f} s",
                                   "pruned": {"value": res["pruned"][0], "ms_per_step": 1e3 * res["pruned"][1],
                                              "what": "the same C restatement with kind-level schema pruning (result-neutral)"},
                                   "build": "gcc -O3 -march=native -fopenmp, state predicate in C, one call set per step"}
        emit(out)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""bench.py - path-instances/s of the hot path on BASELINE.json's headline configuration.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

One step = one pass of the hot path over one incident batch (workloads.Config3): the undirected
<=4-hop Pod->Node enumeration from 100k seed Pods plus the five per-incident instance queries,
with the state predicate fused into emission.  Multi-GPU: graph replicated, one incident batch
per rank (weak scaling), NCCL all-gather of the per-rank instance counts; no data-path collective.

Printed line (rank 0): see the contract in the project brief.  `value` = instances of all ranks /
max-over-ranks device time with inputs resident in HBM; `e2e` = the same through host buffers
(pinned H2D of seeds and timestamps, D2H of every result row inside the timed region).
"""
import argparse
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


def run_reference(args, rank, world):
    """--impl reference: the CPU restatement of the reference's path (oracle/csr_oracle.c, kind
    "port" - Neo4j itself is unavailable on this box), all host threads, bounded sample."""
    if rank != 0:
        return
    from k8s_llm_rca_b200 import workloads
    from oracle.csr import CsrOracle, build
    build()
    w = workloads.Config3(n_objects=args.objects, n_events=args.events)
    orc = CsrOracle(w.snap)
    pods, ev, ts = w.batch(0)
    n = min(len(pods), args.cpu_sample)
    pods, ev, ts = pods[:n], ev[:n], ts[:n]
    threads = os.cpu_count() or orc.max_threads()        # torchrun exports OMP_NUM_THREADS=1: ask for the cores explicitly
    for _ in range(args.warmup):
        w.oracle_run(orc, pods[:max(1, n // 10)], ev[:max(1, n // 10)], ts[:max(1, n // 10)], threads=threads)
    t0 = time.perf_counter()
    inst = 0
    for _ in range(args.steps):
        tot, _ = w.oracle_run(orc, pods, ev, ts, threads=threads)
        inst += tot
    dt = time.perf_counter() - t0
    v = inst / dt
    sample = f"first {n} of {w.n_inc} incidents of the config-3 batch per step"
    emit({
        "metric": METRIC, "value": v, "unit": "instances/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int32", "data": "synthetic", "impl": "reference",
        "config": {"workload": w.name, "objects": args.objects, "incidents_per_step": n},
        "cpu_baseline": {"value": v, "unit": "instances/s", "cores": threads, "kind": "port", "sample": sample},
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
    ap.add_argument("--cpu-sample", type=int, default=50_000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--latency-incidents", type=int, default=10_000,
                    help="config 5: incidents replayed as one batch (drawn with replacement); 0 skips the latency leg")
    ap.add_argument("--latency-singles", type=int, default=200, help="incidents replayed one by one for p50/p99")
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

        def timed_e2e_pipelined(steps):
            """End to end with the copy-out of step k travelling while step k+1 runs: the engine alternates between
            two output arenas (rca_set_double_buffer) and the rows leave through rca_result_copy_async.  Every step
            still uploads its seeds from pinned host memory and downloads all of its rows; ONE pair of events
            brackets all steps (L2 flushes included)."""
            total_words = sum(r.emitted[s] * r.row_len[s] for _, r in res0 for s in range(1, r.levels + 1))
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
                tot, _ = w.run(eng, h_pods.numpy(), h_ev.numpy(), h_ts.numpy(), fetch=fetch_async)
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
        if peer_ok:                                                              # what the last exchange delivered
            last = last_counts[0]
            assert len(last) == world and (last > 0).all(), f"count exchange returned {last}"
        barrier()
        sampler.window(t_win0, time.time())
        clocks = sampler.stop() if rank == 0 else None

    # max over ranks of the device time, sum over ranks of the instances
    red = torch.tensor([ms_res, ms_e2e, ms_serial], dtype=torch.float64, device=dev)
    tot = torch.tensor([inst_res, inst_e2e], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(red, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    ms_res, ms_e2e, ms_serial = red.tolist()
    inst_res, inst_e2e = tot.tolist()

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
        rp.run_batch(mine[:64])                                   # warm-up: message arena upload, metapath cache
        barrier()
        t0 = time.perf_counter()
        out = rp.run_batch(mine)
        torch.cuda.synchronize()
        bt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(bt, op=dist.ReduceOp.MAX)
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
                       "rows_in_batch": int(sum(len(r["rows"]) for r in out)),
                       "what": "per incident: CONTAINS seed lookup + source kind + metapath ladder + instance query + "
                               "message filter + state check per node, LLM answers recorded; single = through "
                               "run_query one statement at a time, batch = one launch set for all incidents"}
        sx.close()
        mx.close()

    if rank == 0:
        peak, peak_src = peaks()
        # dominant kernel = the level launch with the largest share of the step
        (kname, ks), (kms, kbytes, kcalls) = max(per_level.items(), key=lambda kv: kv[1][0])
        achieved = (kbytes / kcalls) / (kms / kcalls * 1e-3) / 1e9
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            try:
                traffic = json.load(open(tp)).get(f"{kname}/level{ks}")
            except Exception:
                traffic = None
        out = {
            "metric": METRIC, "value": inst_res / (ms_res * 1e-3), "unit": "instances/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_res / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": {"workload": w.name, "objects": args.objects, "incidents_per_step_per_gpu": int(len(pods)),
                       "instances_per_step": inst_res // args.steps, "max_hops": w.max_hops,
                       "graph_bytes_in_hbm": int(eng.graph_bytes()),
                       "l2": f"flushed between timed steps (256 MiB write); graph ({eng.graph_bytes() / 1e6:.0f} MB) also exceeds the 126 MB L2",
                       "parallelism": f"graph replicated x{world}, one incident batch per GPU, all-gather of counts"
                                      + (" over peer memory (rca_peer_allgather)" if peer_ok else " (NCCL)" if world > 1 else "")},
            "e2e": {"value": inst_e2e / (ms_e2e * 1e-3), "unit": "instances/s",
                    "h2d_bytes_per_step": int(pods.nbytes + ev.nbytes + 2 * ts.nbytes),     # the five instance queries share one seed list
                    "d2h_bytes_per_step": int(d2h[0] // args.steps), "ms_per_step": ms_e2e / args.steps,
                    "wall_ms_per_step": 1e3 * wall_e2e / args.steps,
                    "how": "every step: seeds host->device from pinned memory, rca_expand_many, ALL rows device->host into "
                           "pinned memory; the copy-out of a step overlaps the next step (two output arenas, "
                           "rca_result_copy_async); one pair of CUDA events around all steps, L2 flushes included",
                    "serial_ms_per_step": ms_serial / args.steps},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "kernel": f"level {ks} of the {kname} plan (its count + scan + fill launches)",
                         "peak_source": peak_src, "bytes_per_launch": kbytes // kcalls,
                         "ms_per_launch": kms / kcalls, "share_of_step": (kms / kcalls) / (ms_res / args.steps)},
            "clocks": clocks,
            "latency": latency,
        }
        if world == 1 and not args.no_cpu_baseline:
            from oracle.csr import CsrOracle, build
            build()
            orc = CsrOracle(w.snap)
            n = min(len(pods), args.cpu_sample)
            cores = os.cpu_count() or orc.max_threads()
            w.oracle_run(orc, pods[:n // 10], ev[:n // 10], ts[:n // 10], threads=cores)
            t0 = time.perf_counter()
            ctot, _ = w.oracle_run(orc, pods[:n], ev[:n], ts[:n], threads=cores)
            dt = time.perf_counter() - t0
            out["cpu_baseline"] = {"value": ctot / dt, "unit": "instances/s", "cores": cores, "kind": "port",
                                   "sample": f"first {n} of {len(pods)} incidents of the same batch, one pass, {dt:.1f} s"}
        emit(out)
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""bench.py — headline benchmark of the batched Filter+Score path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--config C] [--impl ours|reference]

One "step" = one full evaluation of the pending-pod batch against every node of this rank's shard:
reduce pass (Filter + bitmap + normaliser reductions), the cross-shard merges (N > 1), emit pass
(Score / NormalizeScore / weights / sum / selectHost key). The timed region starts with the node and
pod tables resident in HBM. Prints ONE JSON line on rank 0.

  N = 1   workload = config 2 (5k nodes × 8 GPUs, 100k mixed pods, FlexGPU + Fit filter,
          FlexGPU + LeastAllocated scores) — BASELINE.json configs[1]
  N > 1   weak scaling: every rank holds a config-2-sized node shard (5k nodes) of a 5k·N-node
          cluster, all ranks see the same 100k pods; per-pod reductions are merged (MAX / SUM) by
          the engine's peer-memory kernels over NVLink / NVSwitch multicast (--merge nccl: NCCL
          all-reduces instead). `--config 5` runs the 50k-node / 1M-pod config
          node-sharded instead (strong scaling).

--impl reference times the reference's own algorithm on the host cores: the Go plugins cannot be built
here (no Go toolchain), so this is the object-walking C++ restatement under oracle/ ("port").
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "pod_x_node_filter_score_evals_per_sec"
UNIT = "evals/s"
BATCH = 65536  # pods per evaluation batch for configs whose score matrix would not fit otherwise


def workload_name(cfg_id, n_nodes, n_pods, n_gpus, scaling):
    return (f"C{cfg_id}: {n_nodes} nodes x 8 GPUs, {n_pods} pending pods, FlexGPU+NodeResourcesFit filter, "
            f"FlexGPU+{'NodeResourcesAllocatable(Least)' if cfg_id == 1 else 'LeastAllocated'} scores, "
            f"{'single GPU' if n_gpus == 1 else f'nodes sharded over {n_gpus} GPUs ({scaling})'}")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device, uuid=None):
        self.device = device
        self.uuid = uuid
        self.rows = []
        self.source = "nvidia-smi"
        self.window = None  # (t0, t1) in time.perf_counter() terms: the timed region
        self.period_s = float(os.environ.get("BENCH_CLOCK_PERIOD_MS", "4")) * 1e-3  # NVML sampling period
        self._stop = threading.Event()
        self.ready = threading.Event()  # set once the first sample is in
        self._t = None

    def _nvml_handle(self):
        """NVML handle of the CUDA device (by UUID: CUDA_VISIBLE_DEVICES may renumber), or None"""
        try:
            import pynvml
            pynvml.nvmlInit()
            if self.uuid:
                for u in ("GPU-" + self.uuid, self.uuid):
                    try:
                        return pynvml, pynvml.nvmlDeviceGetHandleByUUID(u)
                    except Exception:
                        pass
            if os.environ.get("CUDA_VISIBLE_DEVICES"):
                return None  # the index may be renumbered: leave it to nvidia-smi's view of the same index
            return pynvml, pynvml.nvmlDeviceGetHandleByIndex(self.device)
        except Exception:
            return None

    def _run_nvml(self, nv, h):
        """the same counters as the nvidia-smi query through NVML: a sample every few ms, so that the short timed region
        (tens of ms) is covered by more than one"""
        mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
        bits = (("hw_slowdown", nv.nvmlClocksEventReasonHwSlowdown), ("hw_thermal_slowdown", nv.nvmlClocksEventReasonHwThermalSlowdown),
                ("sw_thermal_slowdown", nv.nvmlClocksEventReasonSwThermalSlowdown), ("sw_power_cap", nv.nvmlClocksEventReasonSwPowerCap))
        while not self._stop.is_set():
            sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
            r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
            self.rows.append([str(self.device), str(sm), str(mx), "", hex(r)] + ["Active" if r & b else "Not Active" for _, b in bits]
                             + [time.perf_counter()])
            self.ready.set()
            self._stop.wait(self.period_s)

    def _run(self):
        got = self._nvml_handle()
        if got is not None:
            try:
                self.source = "nvml"
                self._run_nvml(*got)
                return
            except Exception:
                self.source = "nvidia-smi"
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.device)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")] + [time.perf_counter()])
                    self.ready.set()
            except Exception:
                pass
            self._stop.wait(0.2)

    def start(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join(timeout=6)
        sm, mx, reasons = [], [], set()
        rows = self.rows
        if self.window is not None:  # the samples taken while the device-timed loop ran, if any fell inside
            inside = [r for r in rows if self.window[0] <= r[-1] <= self.window[1]]
            rows = inside or rows
        for r in rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except Exception:
                continue
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            for nm, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "source": self.source}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm), "source": self.source,
                "window": "timed region" if self.window is not None and rows is not self.rows else "warm-up to end of the device-timed loops"}


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def cpu_port_baseline(fgs, synth, cfg, budget_s=12.0, workers=None):
    """Time the oracle (object-walking port of the Go plugins) on the host cores on a bounded sample."""
    from oracle import oracle
    cores = os.cpu_count() or 1
    if workers is None:
        workers = cores
    N, P = synth.n_nodes, synth.n_pending
    probe = min(P, 32)
    t0 = time.perf_counter()
    oracle.evaluate(synth.snapshot, synth.pending, cfg, 0, probe, workers=workers, planes=())
    dt = max(time.perf_counter() - t0, 1e-6)
    n = int(min(P, max(probe, probe * budget_s / dt)))
    t0 = time.perf_counter()
    oracle.evaluate(synth.snapshot, synth.pending, cfg, 0, n, workers=workers, planes=())
    dt = time.perf_counter() - t0
    return {"value": n * N / dt, "unit": UNIT, "cores": cores, "threads": workers, "kind": "port",
            "sample": f"first {n} pending pods x all {N} nodes, {workers} worker threads in the reference's chunked "
                      f"parallelize.Until shape, {dt:.1f} s wall"}


def run_reference(args):
    """--impl reference: the CPU restatement of the Go plugins, all host threads, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import flex_gpu_scheduler_b200 as fgs
    from oracle import oracle
    cfg_id = args.config or 2
    n_nodes = fgs.CONFIGS[cfg_id]["n_nodes"] * (args.gpus if (args.gpus > 1 and cfg_id != 5) else 1)
    cores = os.cpu_count() or 1
    # calibrate the per-step sample so that the whole run stays within a few minutes
    s = fgs.config_synth(cfg_id, n_nodes=n_nodes, n_pending=4096)
    cfg = fgs.config_plugins(cfg_id)
    t0 = time.perf_counter()
    oracle.evaluate(s.snapshot, s.pending, cfg, 0, 16, workers=cores, planes=())
    per_pod = max((time.perf_counter() - t0) / 16, 1e-6)
    total_steps = args.steps + args.warmup
    n = int(max(16, min(4096, (120.0 / total_steps) / per_pod)))
    for _ in range(args.warmup):
        oracle.evaluate(s.snapshot, s.pending, cfg, 0, n, workers=cores, planes=())
    t0 = time.perf_counter()
    for _ in range(args.steps):
        oracle.evaluate(s.snapshot, s.pending, cfg, 0, n, workers=cores, planes=())
    dt = time.perf_counter() - t0
    val = args.steps * n * n_nodes / dt
    scaling = "strong" if cfg_id == 5 else "weak"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "int64",
        "data": "synthetic",
        "config": {"workload": workload_name(cfg_id, n_nodes, fgs.CONFIGS[cfg_id]["n_pending"], args.gpus, scaling),
                   "note": "Go toolchain absent: oracle/ref_walk.cpp, the object-walking C++ restatement of the Go plugins, "
                           "percentageOfNodesToScore=100"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "threads": cores, "kind": "port",
                         "sample": f"each step = first {n} pending pods x all {n_nodes} nodes"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def run_ours(args):
    import torch
    import torch.distributed as dist
    import flex_gpu_scheduler_b200 as fgs
    from flex_gpu_scheduler_b200 import _lib as L
    from flex_gpu_scheduler_b200 import distributed as D

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cfg_id = args.config or 2
    base = fgs.CONFIGS[cfg_id]
    strong = cfg_id == 5
    scaling = "strong" if strong else "weak"
    n_total = base["n_nodes"] * (1 if strong else world)
    n_pods = args.pods or base["n_pending"]
    # every rank generates the same cluster (deterministic) and packs only its node range
    synth = fgs.config_synth(cfg_id, n_nodes=n_total, n_pending=n_pods)
    n0, n1 = D.shard_bounds(n_total, world, rank)
    plugins = fgs.config_plugins(cfg_id)
    stream = torch.cuda.Stream()
    eng = fgs.Engine(device=local, n_scalars=2, node_index_base=n0, total_nodes=n_total, stream=stream.cuda_stream)
    eng.set_plugins(plugins)
    pk_n, pk_p = fgs.Packer(2), fgs.Packer(2)
    eng.upload_nodes(pk_n.nodes(synth.snapshot, n0, n1))
    n_shard = n1 - n0

    # pods go through in batches when the score matrix of the whole set would not fit comfortably
    n_pad = (max(n_shard, 1) + 1023) // 1024 * 1024
    batch = n_pods if n_pods * n_pad * 8 <= 100 * 2 ** 30 else BATCH  # 180 GB of HBM3e per GPU
    batches = [(b, min(n_pods, b + batch)) for b in range(0, n_pods, batch)]
    packers = [fgs.Packer(2) for _ in batches]
    tables = [pk.pods(synth.pending, b0, b1) for pk, (b0, b1) in zip(packers, batches)]
    multi = len(batches) > 1

    # without NodeResourcesAllocatable among the score plugins the emit pass does not read the merged counts: their
    # all-reduce overlaps it (distributed.merge_after_reduce) and is waited for at the end of the step
    defer_counts = not (plugins.score_mask & L.FGS_PL_ALLOCATABLE)

    views = {}  # torch views of the engine's per-pod buffers, made once per buffer (not per step)

    def tensors(view):
        key = (int(view.reduce_max), int(view.p_pad))
        if key not in views:
            views[key] = D.device_views(torch, view)
        return views[key]

    # cross-shard merges: the engine's own kernels over NVLink peer memory (NVSwitch multicast where the box has
    # it), NCCL all-reduces only if symmetric memory cannot be set up (or --merge nccl)
    px, merge_name = None, "none (single GPU)"
    if world > 1:
        merge_name = "NCCL all-reduce"
        if args.merge != "nccl":
            px, why = D.try_peer_exchange(torch, dist, eng, max(b1 - b0 for b0, b1 in batches), multicast=(args.merge != "peer"))
            if px is not None:
                merge_name = ("peer-memory kernel, NVSwitch multicast (multimem.ld_reduce / multimem.st)" if px.multicast
                              else "peer-memory kernel, pushed slices over NVLink peer pointers")
            else:
                merge_name += " (peer memory unavailable: %s)" % why

    def merges_after_reduce(view):
        if px is not None:
            eng.peer_merge_reduce()
        elif world > 1:
            rm, rs, _ = tensors(view)
            return D.merge_after_reduce(dist, rm, rs, defer_counts)
        return None

    def merges_after_emit(view):
        if px is not None:
            eng.peer_merge_select()
        elif world > 1:
            D.merge_after_emit(dist, tensors(view)[2])

    with torch.cuda.stream(stream):
        lib = L.load()
        step_evals = n_pods * n_total  # whole job: all pods x all nodes of all shards

        fused = False

        def one_step(ev=None):
            for t in tables:
                if multi:  # only one batch is resident at a time: restaging is part of the step
                    eng.stage_pods(t)
                v = eng.device_view() if multi else view
                if ev is not None:
                    e = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
                    ev.append(e)
                    e[0].record(stream)
                if fused:  # one engine, nothing between Filter and Score: per-pod prepass, then ONE pass over pod x node
                    eng.pass_fused_prepass()
                else:
                    eng.pass_reduce()
                if ev is not None:
                    e[1].record(stream)
                pending = merges_after_reduce(v)
                if ev is not None:
                    e[2].record(stream)
                if fused:
                    eng.pass_fused_main()
                else:
                    eng.pass_emit()
                if ev is not None:
                    e[3].record(stream)
                merges_after_emit(v)
                if ev is not None:
                    e[4].record(stream)
                if pending is not None:
                    pending.wait()

        eng.stage_pods(tables[0])
        view = eng.device_view()
        eng.synchronize()
        fused = (not args.two_pass) and eng.fused_eligible()
        try:
            dev_uuid = str(torch.cuda.get_device_properties(local).uuid)
        except Exception:
            dev_uuid = None
        sampler = ClockSampler(local, dev_uuid)
        if world > 1 and "BENCH_CLOCK_PERIOD_MS" not in os.environ:
            sampler.period_s = 0.02  # NVML queries share a driver lock with kernel launches; the ranks run in lockstep here
        if rank == 0:
            sampler.start()  # before the warm-up: NVML takes some ms to come up, the timed region is short
        for _ in range(args.warmup):
            one_step()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        if rank == 0:
            sampler.ready.wait(2.0)
        launches0 = eng.timings()["kernel_launches"]
        evs = []
        e_start, e_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        t_w0 = time.perf_counter()
        e_start.record(stream)
        for k in range(args.steps):
            one_step(evs)
        e_end.record(stream)
        torch.cuda.synchronize()
        sampler.window = (t_w0, time.perf_counter())
        if world > 1:
            dist.barrier()
        if px is not None:
            eng.peer_status()  # raises if a merge gave up waiting for a peer
        ms = e_start.elapsed_time(e_end)
        launches = eng.timings()["kernel_launches"] - launches0
        reduce_ms = sum(e[0].elapsed_time(e[1]) for e in evs) / args.steps
        emit_ms = sum(e[2].elapsed_time(e[3]) for e in evs) / args.steps
        # medians: a merge also waits for the slowest rank's previous pass, so one host hiccup on any rank lands here
        merge1_ms = sorted(e[1].elapsed_time(e[2]) for e in evs)[len(evs) // 2]
        merge2_ms = sorted(e[3].elapsed_time(e[4]) for e in evs)[len(evs) // 2]
        if world > 1:
            t = torch.tensor([ms, reduce_ms, emit_ms, merge1_ms, merge2_ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms, reduce_ms, emit_ms, merge1_ms, merge2_ms = [float(x) for x in t.tolist()]

        # ---- end to end through the public C-ABI call with HOST buffers ---------------------------------
        # fgs_evaluate: H2D of the pod table from pinned staging + both passes + D2H of the per-pod results.
        # With N > 1 the staged calls are used so that the NCCL merges sit between the passes.
        res_holder = [None]

        def e2e_step():
            for t in tables:
                if world == 1:
                    res_holder[0] = eng.evaluate(t)
                else:
                    eng.stage_pods(t)
                    v = eng.device_view()
                    eng.pass_reduce()
                    pending = merges_after_reduce(v)
                    eng.pass_emit()
                    merges_after_emit(v)
                    if pending is not None:
                        pending.wait()
                    res_holder[0] = eng.fetch_results()

        e2e_steps = max(1, min(args.steps, 5))
        e2e_step()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        e_s, e_e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e_s.record(stream)
        for _ in range(e2e_steps):
            e2e_step()
        e_e.record(stream)
        torch.cuda.synchronize()
        e2e_wall = time.perf_counter() - t0
        e2e_ms = max(e_s.elapsed_time(e_e), 1e3 * e2e_wall)  # host-side packing into pinned memory counts too
        if world > 1:
            t = torch.tensor([e2e_ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e2e_ms = float(t.item())
        # The same through fgs_submit / fgs_collect: a stream of batches with two in flight — every step still
        # stages its pod table on the host, copies it H2D and reads its per-pod results back D2H, but the host side of
        # step k+1 and the copy run while the kernels of step k do.
        # With N > 1 (peer-memory merges) the same calls include the cross-shard merges, on their own stream.
        pipe_ms, pipe_steps = None, 0
        if (world == 1 or px is not None) and not multi:
            pipe_steps = max(10, args.steps)
            depth = 2 if world == 1 else 3  # batches in flight (the sharded pipeline defers a batch's emit pass by one submit)

            def pipe_loop(k, sub):
                inflight = 0
                for _ in range(k):
                    sub()
                    inflight += 1
                    if inflight == depth:
                        res_holder[0] = eng.collect()
                        inflight -= 1
                for _ in range(inflight):
                    res_holder[0] = eng.collect()

            pipe_loop(depth, lambda: eng.submit(tables[0]))  # every pipeline slot now holds the batch
        # N > 1, tables resident in HBM (both pipeline slots hold the batch after the loop above): the headline loop
        # again with two batches in flight, so that every merge overlaps a pass of the other batch. Each step still runs
        # both passes, both merges, the per-pod epilogue and the D2H of the per-pod results.
        ov_ms = None
        if px is not None and not multi:
            def ov_loop(k):
                pipe_loop(k, lambda: eng.resubmit(n_pods))
            ov_loop(max(3, args.warmup))
            torch.cuda.synchronize()
            dist.barrier()
            launches0 = eng.timings()["kernel_launches"]
            o_s, o_e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            o_s.record(stream)
            ov_loop(args.steps)
            o_e.record(stream)  # after the last collect returned: every stream of the pipeline has drained
            torch.cuda.synchronize()
            sampler.window = (sampler.window[0], time.perf_counter())  # both device-timed loops
            dist.barrier()
            eng.peer_status()
            ov_launches = eng.timings()["kernel_launches"] - launches0
            t = torch.tensor([o_s.elapsed_time(o_e)], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ov_ms = float(t.item())
        # the clock sampler (an nvidia-smi process every 200 ms) has covered the device-timed loops; it stops here: its
        # driver queries delay kernel launches, which the host-paced pipeline below would feel
        clocks = sampler.stop() if rank == 0 else None
        if (world == 1 or px is not None) and not multi:
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            t0 = time.perf_counter()
            pipe_loop(pipe_steps, lambda: eng.submit(tables[0]))
            torch.cuda.synchronize()
            pipe_ms = 1e3 * (time.perf_counter() - t0)
            if world > 1:
                t = torch.tensor([pipe_ms], device="cuda", dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                pipe_ms = float(t.item())

    # ---- true scheduling cycles: one pod at a time with assume + Reserve between pods ----------
    # A second engine on the same cluster (the sequence rewrites node rows); a bounded prefix of the queue.
    seq = None
    if world == 1 and not args.no_sequential:
        n_seq = min(n_pods, args.seq_pods)
        eng_s = fgs.Engine(device=local, n_scalars=2)
        eng_s.set_plugins(plugins)
        pk_s = fgs.Packer(2)
        seq_table = pk_s.pods(synth.pending, 0, n_seq)
        best_ms, wall_ms, sres = None, None, None
        for _ in range(2):  # the first call pays the buffer allocations
            eng_s.upload_nodes(pk_n.nodes(synth.snapshot, n0, n1))
            t0 = time.perf_counter()
            sres = eng_s.schedule_sequential(seq_table)
            wall_ms = 1e3 * (time.perf_counter() - t0)
            best_ms = eng_s.timings()["seq_ms"]
        ts = eng_s.timings()
        seq = {"cycles_per_sec": n_seq / (wall_ms * 1e-3), "kernel_cycles_per_sec": n_seq / (best_ms * 1e-3), "pods": n_seq,
               "committed": int(sres.committed.sum()), "unschedulable": int((sres.feasible_count == 0).sum()),
               "ctas": abs(ts["seq_ctas"]), "exchange": "dsmem cluster" if ts["seq_ctas"] < 0 else "global mailboxes",
               "api": "fgs_schedule_sequential (host pod table in; Filter+Score+selectHost, assume, FlexGPU.Reserve per pod; "
                      "per-pod node / device index out)"}
        eng_s.close()

    if rank == 0:
        res = res_holder[0]
        P, N = n_pods, n_shard
        p_pad = sum((b1 - b0 + 63) // 64 * 64 for b0, b1 in batches)
        S = 2
        # algorithmic bytes of the dominant kernel (emit pass): one int64 score per (pod,node) + each table once
        nd_cols, pd_cols = 15 + 2 * S, 8 + S
        emit_bytes = P * N * 8 + N * nd_cols * 8 + P * (pd_cols + 4) * 8
        path_bytes = P * N * (8 + 1 / 8) + 2 * (N * nd_cols * 8 + P * pd_cols * 8) + P * 10 * 8
        if fused:  # the one pass writes the score AND the feasibility bit of every (pod,node): SURVEY.md §8d's B_alg
            emit_bytes = int(P * N * (8 + 1 / 8)) + N * nd_cols * 8 + P * (pd_cols + 4) * 8
            path_bytes = emit_bytes + P * pd_cols * 8
        peak, peak_src = measured_peak_gbs()
        achieved = emit_bytes / (emit_ms * 1e-3) / 1e9  # emit_ms sums the step's emit launches (one per pod batch)
        traffic = None
        tp = os.path.join(ROOT, "profiles", "r01_fused_traffic.json" if fused else "r01_emit_traffic.json")
        if os.path.exists(tp) and cfg_id == 2 and len(batches) == 1 and n_pods == base["n_pending"]:
            try:
                traffic = json.load(open(tp)).get("dram_bytes_per_launch")
            except Exception:
                traffic = None
        sync_ms_per_step = ms / args.steps
        if ov_ms is not None:  # the headline for N > 1: two batches in flight, merges overlapped
            ms, launches = ov_ms, ov_launches
        value = args.steps * step_evals / (ms * 1e-3)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "int64",
            "data": "synthetic",
            "config": {"workload": workload_name(cfg_id, n_total, n_pods, world, scaling), "nodes_per_gpu": n_shard,
                       "l2": "outputs (8 B per eval, %.1f GB per step per GPU) exceed the 126 MB L2; node and pod tables "
                             "(%.1f MB) are L2-resident by design" % (P * N * 8 / 1e9, (N * nd_cols * 8 + P * pd_cols * 8) / 1e6),
                       "parity": "all nodes evaluated for every pod (percentageOfNodesToScore=100)",
                       "pod_batches": len(batches), "merge": merge_name},
            "kernel_ms": dict({"flexmax_prepass": reduce_ms, "fused_pass": emit_ms} if fused else {"reduce_pass": reduce_ms, "emit_pass": emit_ms},
                              **({"merge_after_reduce_median": merge1_ms, "merge_after_emit_median": merge2_ms, "one_batch_at_a_time_ms_per_step": sync_ms_per_step,
                                  "note": "kernel_ms: measured with one batch at a time (pass, merge, pass, merge in sequence); "
                                          "ms_per_step / value: " + ("up to three batches in flight (fgs_resubmit / fgs_collect), merges overlap the other "
                                                                     "batch's pass" if ov_ms is not None else "the same sequence")} if world > 1 else {})),
            "scheduling_cycles_per_sec": args.steps * n_pods / (ms * 1e-3),
            "scheduling_cycles_note": "batched: every pod against the same snapshot, no commit between pods; "
                                      "'sequential' below commits each placement before the next pod",
            "sequential": seq,
            "roofline": {"bound": "hbm", "kernel": ("emit_narrow<SPEC, fused> (Filter + Score + NormalizeScore + selectHost key in one pass, int32)" if fused else
                                                    "emit_narrow (int32 emit pass)" if eng.timings()["emit_int32"] else "eval_kernel<EMIT> (int64 emit pass)"), "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": emit_bytes // len(batches), "launches_per_step": len(batches),
                         "path_frac": (path_bytes / ((reduce_ms + emit_ms) * 1e-3) / 1e9) / peak},
            "e2e": ({"value": pipe_steps * step_evals / (pipe_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": pd_cols * p_pad * 8,
                     "d2h_bytes_per_step": 4 * p_pad * 8, "steps": pipe_steps,
                     "api": "fgs_submit + fgs_collect, " + ("two" if world == 1 else "three") + " batches in flight (host pod table in, per-pod selected node / count / "
                            "error out; every step stages, copies H2D and reads back D2H"
                            + (")" if world == 1 else "; cross-shard merges inside the calls, on their own stream)"),
                     "single_call": {"value": e2e_steps * step_evals / (e2e_ms * 1e-3), "steps": e2e_steps,
                                     "api": "fgs_evaluate, one blocking call per step" if world == 1 else
                                            "fgs_stage_pods + fgs_pass_reduce + merge + fgs_pass_emit + merge + fgs_fetch_results"}}
                    if pipe_ms else
                    {"value": e2e_steps * step_evals / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": pd_cols * p_pad * 8,
                     "d2h_bytes_per_step": 4 * p_pad * 8, "steps": e2e_steps,
                     "api": "fgs_evaluate (host pod table in, per-pod selected node / count / error out)"
                     if world == 1 else "fgs_stage_pods + fgs_pass_reduce + merge + fgs_pass_emit + merge + fgs_fetch_results (merge: see config)"}),
            "gpu_launches": int(launches),
            "clocks": clocks,
            "selected_nodes_checksum": int(res.selected_node.astype("int64").sum()) if res is not None else None,
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_port_baseline(fgs, synth, plugins)
        print(json.dumps(line))
    eng.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", type=int, default=0, help="1..5 (default 2)")
    ap.add_argument("--pods", type=int, default=0, help="override the number of pending pods")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sequential", action="store_true")
    ap.add_argument("--two-pass", action="store_true", help="N = 1: reduce pass + emit pass even where Filter and Score can run in one pass")
    ap.add_argument("--merge", default="auto", choices=["auto", "peer", "nccl"],
                    help="N > 1: cross-shard merges through the engine's peer-memory kernels (auto: NVSwitch multicast if there is one; "
                         "peer: plain peer pointers) or through NCCL all-reduces")
    ap.add_argument("--seq-pods", type=int, default=20000, help="pods of the sequential (commit between pods) measurement")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())

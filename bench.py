#!/usr/bin/env python
"""bench.py — headline benchmark of the batched Filter+Score path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--config C] [--impl ours|reference]

One "step" = one full evaluation of the WHOLE pending-pod set of the configuration against every node: per pod batch
the Filter pass (bitmap + normaliser reductions), the cross-shard merges (N > 1), the emit pass (Score / NormalizeScore
/ weights / sum / selectHost key), the per-pod epilogue and the D2H of the per-pod results. The timed region of `value`
starts with the node and pod tables resident in HBM; `e2e` starts from pinned host tables. ONE JSON line on rank 0.

  N = 1   workload = config 2 (5 000 nodes x 8 GPUs, 100 000 mixed pods, FlexGPU + Fit filter, FlexGPU + LeastAllocated
          scores: BASELINE.json configs[1], the configuration the metric is quoted on for one B200); the line also
          carries `c5_single_gpu` = config 5 on the same GPU, the denominator of the strong-scaling curve
  N > 1   workload = config 5 (50 000 nodes, 1 000 000 pods in 65 536-pod batches) with the NODES sharded over the N GPUs
          (strong scaling: BASELINE.json configs[4], north_star's scaling target), all ranks see the same pods, per-pod
          reductions merged per batch by the engine's peer-memory kernels over NVLink / NVSwitch multicast (--merge nccl:
          NCCL all-reduces instead); `c2_weak` = the round-1 weak-scaling workload (a config-2 shard per GPU) beside it
  --config C forces configuration C at any N (1, 2: one batch; 4: 20 000 nodes; 5: as above); --config 3 is the
          Coscheduling line (PodGroup PreFilter checks per second)

The SAME loop runs at every N: the batches of the job are resident in the engine's pipeline slots and fgs_resubmit /
fgs_collect walk them with two (N = 1) or three (N > 1: a merge overlaps the other batch's pass) batches in flight.

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


def default_config(n_gpus, forced):
    """N = 1 → config 2 (BASELINE's one-B200 configuration); N > 1 → config 5 node-sharded (north_star's scaling target)"""
    return forced or (2 if n_gpus == 1 else 5)


def cpu_port_baseline(fgs, synth, cfg, budget_s=10.0):
    """Time the oracle (object-walking port of the Go plugins) on the host cores on a bounded sample: once with the
    reference's default of 16 workers (parallelize/parallelism.go:27), once with every core of the box."""
    from oracle import oracle
    cores = os.cpu_count() or 1
    N, P = synth.n_nodes, synth.n_pending

    def run(workers):
        probe = min(P, 32)
        t0 = time.perf_counter()
        oracle.evaluate(synth.snapshot, synth.pending, cfg, 0, probe, workers=workers, planes=())
        dt = max(time.perf_counter() - t0, 1e-6)
        n = int(min(P, max(probe, probe * budget_s / dt)))
        t0 = time.perf_counter()
        oracle.evaluate(synth.snapshot, synth.pending, cfg, 0, n, workers=workers, planes=())
        dt = time.perf_counter() - t0
        return n * N / dt, n, dt

    v16, n16, dt16 = run(16)
    line = {"value": v16, "unit": UNIT, "cores": cores, "threads": 16, "kind": "port",
            "sample": f"first {n16} pending pods x all {N} nodes, 16 worker threads (the reference's default) in its chunked "
                      f"parallelize.Until shape, {dt16:.1f} s wall"}
    if cores != 16:
        va, na, dta = run(cores)
        line["all_cores"] = {"value": va, "threads": cores, "sample": f"first {na} pods x all {N} nodes, {dta:.1f} s wall"}
    else:
        line["all_cores"] = {"value": v16, "threads": 16, "sample": "the box has 16 cores: same run"}
    return line


def run_reference(args):
    """--impl reference: the CPU restatement of the Go plugins, all host threads, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import flex_gpu_scheduler_b200 as fgs
    from oracle import oracle
    cfg_id = default_config(args.gpus, args.config)
    if cfg_id == 3:
        return run_config3(args, reference=True)
    if cfg_id == 4:
        return run_config4(args, reference=True)
    n_nodes = fgs.CONFIGS[cfg_id]["n_nodes"]
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
    scaling = "strong"  # as the GPU arm labels it: the node count of the configuration is fixed, the GPUs share it
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "int64",
        "data": "synthetic",
        "config": {"workload": workload_name(cfg_id, n_nodes, fgs.CONFIGS[cfg_id]["n_pending"], args.gpus, scaling),
                   "note": "Go toolchain absent: oracle/ref_walk.cpp, the object-walking C++ restatement of the Go plugins, "
                           "percentageOfNodesToScore=100; a CPU run: the GPU count only selects the configuration"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "threads": cores, "kind": "port",
                         "sample": f"each step = first {n} pending pods x all {n_nodes} nodes"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def emit_traffic_record():
    """dram bytes per launch of the emit pass from the committed ncu capture — only while the kernel source is the one
    that was profiled (sha256 of eval_narrow.cu recorded beside the number), else null"""
    import hashlib
    for name in ("r02_emit_traffic.json", "r01_emit_traffic.json"):
        tp = os.path.join(ROOT, "profiles", name)
        if not os.path.exists(tp):
            continue
        try:
            rec = json.load(open(tp))
            src = os.path.join(ROOT, "flex-gpu-scheduler_b200", "csrc", "eval_narrow.cu")
            sha = hashlib.sha256(open(src, "rb").read()).hexdigest()[:16]
            if rec.get("kernel_src_sha16") == sha:
                return rec.get("dram_bytes_per_launch"), "profiles/" + name
            return None, f"profiles/{name} is older than eval_narrow.cu (sha {sha}): not quoted"
        except Exception:
            continue
    return None, "no ncu capture committed"


class Workload:
    """One configuration on this process's GPU (its node shard when world > 1): engine, resident batches, loops."""

    def __init__(self, torch, dist, fgs, D, L, args, cfg_id, n_total, n_pods, world, rank, local, stream):
        self.torch, self.dist, self.fgs, self.D, self.L = torch, dist, fgs, D, L
        self.cfg_id, self.n_total, self.n_pods, self.world, self.rank = cfg_id, n_total, n_pods, world, rank
        self.stream = stream
        self.synth = fgs.config_synth(cfg_id, n_nodes=n_total, n_pending=n_pods)
        self.n0, self.n1 = D.shard_bounds(n_total, world, rank)
        self.n_shard = self.n1 - self.n0
        self.plugins = fgs.config_plugins(cfg_id)
        self.eng = fgs.Engine(device=local, n_scalars=2, node_index_base=self.n0, total_nodes=n_total, stream=stream.cuda_stream)
        self.eng.set_plugins(self.plugins)
        self.pk_n = fgs.Packer(2)
        self.eng.upload_nodes(self.pk_n.nodes(self.synth.snapshot, self.n0, self.n1))
        # pods go through in batches when the score matrix of the whole set would not fit comfortably
        n_pad = (max(self.n_shard, 1) + 1023) // 1024 * 1024
        n_pad_single = (n_total + 1023) // 1024 * 1024
        one = n_pods * n_pad_single * 8 <= 100 * 2 ** 30  # 180 GB of HBM3e per GPU; the same batching at every N
        batch = n_pods if one else BATCH
        self.batches = [(b, min(n_pods, b + batch)) for b in range(0, n_pods, batch)]
        # pinned packers: the engine copies the columns straight from these tables by DMA (fgs_packer_pin)
        self.packers = [fgs.Packer(2, pin=True) for _ in self.batches]
        self.tables = [pk.pods(self.synth.pending, b0, b1) for pk, (b0, b1) in zip(self.packers, self.batches)]
        self.sizes = [b1 - b0 for b0, b1 in self.batches]
        self.px, self.merge_name = None, "none (single GPU)"
        if world > 1:
            self.merge_name = "NCCL all-reduce"
            if args.merge != "nccl":
                self.px, why = D.try_peer_exchange(torch, dist, self.eng, max(self.sizes), multicast=(args.merge != "peer"))
                if self.px is not None:
                    self.merge_name = ("peer-memory kernel, NVSwitch multicast (multimem.ld_reduce / multimem.st)" if self.px.multicast
                                       else "peer-memory kernel, pushed slices over NVLink peer pointers")
                else:
                    self.merge_name += " (peer memory unavailable: %s)" % why
        base = 2 if world == 1 else 3
        self.pipelined = world == 1 or self.px is not None
        nb = len(self.batches)
        # a ring of one slot more than the default lets one more batch be in flight: its H2D copy is issued a pass earlier
        self.ring = nb * -(-(base + 1) // nb)  # a multiple of the job's batches
        self.depth = base + 1 if self.ring > base else base  # batches in flight
        self.views = {}
        self.last_results = [None] * nb

    # -- plumbing ------------------------------------------------------------------------------------------------------
    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()

    def max_over_ranks(self, *vals):
        if self.world == 1:
            return [float(v) for v in vals]
        t = self.torch.tensor(list(vals), device="cuda", dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(x) for x in t.tolist()]

    def _tensors(self, view):
        key = (int(view.reduce_max), int(view.p_pad))
        if key not in self.views:
            self.views[key] = self.D.device_views(self.torch, view)
        return self.views[key]

    # -- one batch at a time, every stage bracketed by events: where the time goes -------------------------------------
    def staged_loop(self, steps):
        torch, eng = self.torch, self.eng
        defer_counts = not (self.plugins.score_mask & self.L.FGS_PL_ALLOCATABLE)
        evs = []
        for _ in range(steps):
            for t in self.tables:
                eng.stage_pods(t)
                v = eng.device_view()
                e = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
                evs.append(e)
                e[0].record(self.stream)
                eng.pass_reduce()
                e[1].record(self.stream)
                pending = None
                if self.px is not None:
                    eng.peer_merge_reduce()
                elif self.world > 1:
                    rm, rs, _ = self._tensors(v)
                    pending = self.D.merge_after_reduce(self.dist, rm, rs, defer_counts)
                e[2].record(self.stream)
                eng.pass_emit()
                e[3].record(self.stream)
                if self.px is not None:
                    eng.peer_merge_select()
                elif self.world > 1:
                    self.D.merge_after_emit(self.dist, self._tensors(v)[2])
                e[4].record(self.stream)
                if pending is not None:
                    pending.wait()
        torch.cuda.synchronize()
        nb = len(self.tables)
        red = sum(e[0].elapsed_time(e[1]) for e in evs) / steps
        emit = sum(e[2].elapsed_time(e[3]) for e in evs) / steps
        m1 = sorted(e[1].elapsed_time(e[2]) for e in evs)[len(evs) // 2]
        m2 = sorted(e[3].elapsed_time(e[4]) for e in evs)[len(evs) // 2]
        red, emit, m1, m2 = self.max_over_ranks(red, emit, m1, m2)
        return {"reduce_ms": red, "emit_ms": emit, "merge1_ms": m1, "merge2_ms": m2, "batches": nb}

    # -- the job through the pipeline ----------------------------------------------------------------------------------
    def _walk(self, steps, sub, keep=False):
        """steps x (every batch of the job once), `depth` batches in flight"""
        eng, nb = self.eng, len(self.tables)
        inflight, order = 0, []
        for _ in range(steps):
            for b in range(nb):
                sub(b)
                order.append(b)
                inflight += 1
                if inflight == self.depth:
                    r = eng.collect()
                    if keep:
                        self.last_results[order[-inflight]] = r
                    inflight -= 1
        while inflight:
            r = eng.collect()
            if keep:
                self.last_results[order[-inflight]] = r
            inflight -= 1

    def fill_slots(self):
        """every pipeline slot gets its batch resident (ring = a multiple of the job's batches)"""
        if self.ring > (2 if self.world == 1 else 3):
            self.eng.set_resident_batches(self.ring)
            assert self.eng.max_in_flight() == self.depth
        eng, nb = self.eng, len(self.tables)
        inflight = 0
        for i in range(self.ring):
            eng.submit(self.tables[i % nb])
            inflight += 1
            if inflight == self.depth:
                eng.collect()
                inflight -= 1
        while inflight:
            eng.collect()
            inflight -= 1

    def resident_loop(self, steps):
        """device-timed: tables resident in HBM, fgs_resubmit / fgs_collect"""
        torch = self.torch
        e_s, e_e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = self.eng.timings()["kernel_launches"]
        e_s.record(self.stream)
        self._walk(steps, lambda b: self.eng.resubmit(self.sizes[b]))
        e_e.record(self.stream)  # after the last collect returned: every stream of the pipeline has drained
        torch.cuda.synchronize()
        return e_s.elapsed_time(e_e), self.eng.timings()["kernel_launches"] - l0

    def host_loop(self, steps, keep=False):
        """end to end: pinned host pod tables in (H2D inside), per-pod results out (D2H inside), fgs_submit / fgs_collect"""
        torch = self.torch
        e_s, e_e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e_s.record(self.stream)
        self._walk(steps, lambda b: self.eng.submit(self.tables[b]), keep=keep)
        e_e.record(self.stream)
        torch.cuda.synchronize()
        wall = 1e3 * (time.perf_counter() - t0)
        return max(e_s.elapsed_time(e_e), wall)

    def blocking_loop(self, steps):
        """the same without a pipeline: one blocking call per batch (staged calls + merges when sharded)"""
        torch, eng = self.torch, self.eng
        defer_counts = not (self.plugins.score_mask & self.L.FGS_PL_ALLOCATABLE)
        t0 = time.perf_counter()
        for _ in range(steps):
            for t in self.tables:
                if self.world == 1:
                    eng.evaluate(t)
                else:
                    self.D.evaluate_sharded(torch, self.dist, eng, t, defer_counts=defer_counts, peer=self.px)
        torch.cuda.synchronize()
        return 1e3 * (time.perf_counter() - t0)

    def parity_record(self, blocks=16):
        """rank 0, outside every timed region: hashed blocks of 16 pods of the job re-evaluated by the oracle on the WHOLE
        cluster and compared with what the (sharded) pipeline returned: count, error bits, selectHost's score, the chosen
        node inside the oracle's max-score tie set"""
        if self.rank != 0:
            return None
        from oracle import oracle
        import numpy as np
        P = self.n_pods
        nblk = max(1, P // 16)
        order = sorted(range(nblk), key=lambda b: ((b + self.cfg_id) * 0x9E3779B1) & 0xffffffff)[:blocks]
        checked, bad = 0, []
        for b in sorted(order):
            p0, p1 = 16 * b, min(P, 16 * b + 16)
            bi = next(i for i, (a, z) in enumerate(self.batches) if a <= p0 < z)
            res = self.last_results[bi]
            if res is None:
                continue
            base = self.batches[bi][0]
            ref = oracle.evaluate(self.synth.snapshot, self.synth.pending, self.plugins, p0, p1, workers=os.cpu_count() or 1)
            sl = slice(p0 - base, p1 - base)
            ok = (ref.pod_err == 0) & (ref.feasible_count >= 1)
            idx = np.nonzero(ok)[0]
            sel = res.selected_node[sl]
            good = (np.array_equal(res.feasible_count[sl], ref.feasible_count) and np.array_equal(res.pod_err[sl], ref.pod_err)
                    and np.array_equal(res.selected_score[sl][idx], ref.max_total[idx]) and np.all(sel[ref.feasible_count == 0] == -1)
                    and np.all(sel[idx] >= 0) and np.array_equal(ref.total[idx, sel[idx]], ref.max_total[idx])
                    and np.all(ref.status[idx, sel[idx]] == 0))
            checked += p1 - p0
            if not good:
                bad.append(p0)
        return {"pods_checked": checked, "nodes": self.n_total, "pass": not bad, "failed_blocks_at": bad,
                "what": "feasible count, pod_err, selectHost score and tie-set membership of the chosen node, against oracle.evaluate on the whole cluster"}

    def close(self):
        if self.px is not None:
            self.px.close()
        self.eng.close()


def measure(W, args, steps, warmup, headline):
    """all loops of one workload; returns the record (device-timed value, kernel_ms, e2e, …) — identical on every rank"""
    eng = W.eng
    nb = len(W.tables)
    step_evals = W.n_pods * W.n_total
    rec = {}
    with W.torch.cuda.stream(W.stream):
        W.staged_loop(2)  # warm: clocks up, level tables built, buffers allocated
        km = W.staged_loop(max(1, min(steps, 5)))
        rec["staged"] = km
        rec["filter_lookup"] = bool(eng.timings()["reduce_lookup"])
        rec["emit_int32"] = bool(eng.timings()["emit_int32"])
        if W.pipelined:
            W.fill_slots()
            W.resident_loop(max(1, warmup))
            W.barrier()
            if headline is not None:
                headline["t0"] = time.perf_counter()
            ms, launches = W.resident_loop(steps)
            if headline is not None:
                headline["t1"] = time.perf_counter()
            W.barrier()
            (ms,) = W.max_over_ranks(ms)
            rec["loop"] = "fgs_resubmit / fgs_collect, %d batches in flight, %d pipeline slots" % (W.depth, W.ring)
        else:  # NCCL merges: the staged sequence is the loop
            W.barrier()
            e_s, e_e = W.torch.cuda.Event(enable_timing=True), W.torch.cuda.Event(enable_timing=True)
            l0 = eng.timings()["kernel_launches"]
            if headline is not None:
                headline["t0"] = time.perf_counter()
            e_s.record(W.stream)
            W.staged_loop(steps)
            e_e.record(W.stream)
            W.torch.cuda.synchronize()
            if headline is not None:
                headline["t1"] = time.perf_counter()
            ms, launches = e_s.elapsed_time(e_e), eng.timings()["kernel_launches"] - l0
            (ms,) = W.max_over_ranks(ms)
            rec["loop"] = "fgs_stage_pods + fgs_pass_reduce + all-reduce + fgs_pass_emit + all-reduce, one batch at a time"
        if W.px is not None:
            eng.peer_status()  # raises if a merge gave up waiting for a peer
        rec.update({"ms": ms, "steps": steps, "value": steps * step_evals / (ms * 1e-3), "launches": int(launches), "batches": nb})
    return rec


def measure_e2e(W, args, steps):
    step_evals = W.n_pods * W.n_total
    rec = {}
    with W.torch.cuda.stream(W.stream):
        if W.pipelined:
            W.host_loop(1)
            W.barrier()
            ms = W.host_loop(steps, keep=True)
            (ms,) = W.max_over_ranks(ms)
            rec["pipe"] = {"ms": ms, "steps": steps, "value": steps * step_evals / (ms * 1e-3)}
        bsteps = max(1, min(steps, 3))
        W.blocking_loop(1)
        W.barrier()
        ms = W.blocking_loop(bsteps)
        (ms,) = W.max_over_ranks(ms)
        rec["blocking"] = {"ms": ms, "steps": bsteps, "value": bsteps * step_evals / (ms * 1e-3)}
        W.barrier()
    return rec


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
    cfg_id = default_config(world, args.config)
    if cfg_id == 3:
        return run_config3(args, reference=False)
    if cfg_id == 4:
        return run_config4(args, reference=False)
    base = fgs.CONFIGS[cfg_id]
    strong = cfg_id == 5 or world == 1 or bool(args.config)
    weak = world > 1 and not strong
    scaling = "weak" if weak else "strong"
    n_total = base["n_nodes"] * (world if weak else 1)
    n_pods = args.pods or base["n_pending"]
    stream = torch.cuda.Stream()
    W = Workload(torch, dist, fgs, D, L, args, cfg_id, n_total, n_pods, world, rank, local, stream)
    try:
        dev_uuid = str(torch.cuda.get_device_properties(local).uuid)
    except Exception:
        dev_uuid = None
    sampler = ClockSampler(local, dev_uuid)
    if world > 1 and "BENCH_CLOCK_PERIOD_MS" not in os.environ:
        sampler.period_s = 0.02  # NVML queries share a driver lock with kernel launches; the ranks run in lockstep here
    if rank == 0:
        sampler.start()  # before the warm-up: NVML takes some ms to come up
        sampler.ready.wait(2.0)
    hl = {}
    rec = measure(W, args, args.steps, args.warmup, hl)
    sampler.window = (hl["t0"], hl["t1"])
    clocks = sampler.stop() if rank == 0 else None  # its driver queries delay launches: it stops before the host-paced loops
    e2e = measure_e2e(W, args, max(args.steps, 10) if len(W.tables) == 1 else args.steps)
    parity = W.parity_record() if not args.no_parity else None

    # ---- extra records ------------------------------------------------------------------------------------------------
    extra = {}
    if world == 1 and cfg_id == 2 and not args.config and not args.no_c5:
        # the denominator of the strong-scaling curve: config 5 on this one GPU, same loops
        W5 = Workload(torch, dist, fgs, D, L, args, 5, fgs.CONFIGS[5]["n_nodes"], fgs.CONFIGS[5]["n_pending"], 1, 0, local, stream)
        r5 = measure(W5, args, max(1, min(args.steps, 3)), 1, None)
        e5 = measure_e2e(W5, args, 2)
        p5 = W5.parity_record(blocks=8) if not args.no_parity else None
        extra["c5_single_gpu"] = {
            "workload": workload_name(5, W5.n_total, W5.n_pods, 1, "strong"), "value": r5["value"], "unit": UNIT, "ms_per_step": r5["ms"] / r5["steps"],
            "steps": r5["steps"], "pod_batches": r5["batches"], "loop": r5["loop"],
            "kernel_ms_per_step": {"filter_pass": r5["staged"]["reduce_ms"], "emit_pass": r5["staged"]["emit_ms"]},
            "e2e": {"value": e5["pipe"]["value"], "ms_per_step": e5["pipe"]["ms"] / e5["pipe"]["steps"]}, "parity": p5,
            "note": "strong-scaling efficiency at N GPUs = value(N) / (N x this value)"}
        W5.close()
    if world > 1 and cfg_id == 5 and not args.config and not args.no_weak:
        # round 1's workload beside it: every GPU keeps a config-2-sized shard (weak scaling)
        W2 = Workload(torch, dist, fgs, D, L, args, 2, fgs.CONFIGS[2]["n_nodes"] * world, fgs.CONFIGS[2]["n_pending"], world, rank, local, stream)
        r2 = measure(W2, args, max(1, min(args.steps, 10)), 3, None)
        e2 = measure_e2e(W2, args, 5)
        extra["c2_weak"] = {"workload": workload_name(2, W2.n_total, W2.n_pods, world, "weak"), "value": r2["value"], "unit": UNIT,
                            "ms_per_step": r2["ms"] / r2["steps"], "steps": r2["steps"], "loop": r2["loop"],
                            "e2e": {"value": e2["pipe"]["value"]} if "pipe" in e2 else {"value": e2["blocking"]["value"]},
                            "kernel_ms_per_step": {"filter_pass": r2["staged"]["reduce_ms"], "emit_pass": r2["staged"]["emit_ms"],
                                                   "merge_after_filter_median": r2["staged"]["merge1_ms"],
                                                   "merge_after_emit_median": r2["staged"]["merge2_ms"]}}
        W2.close()

    # ---- true scheduling cycles: one pod at a time with assume + Reserve between pods ----------
    seq = None
    if world == 1 and not args.no_sequential:
        n_seq = min(W.n_pods, args.seq_pods)
        eng_s = fgs.Engine(device=local, n_scalars=2)
        eng_s.set_plugins(W.plugins)
        pk_s = fgs.Packer(2)
        seq_table = pk_s.pods(W.synth.pending, 0, n_seq)
        best_ms, wall_ms, sres = None, None, None
        for _ in range(2):  # the first call pays the buffer allocations
            eng_s.upload_nodes(W.pk_n.nodes(W.synth.snapshot, W.n0, W.n1))
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
        P, N = W.n_pods, W.n_shard
        nb = len(W.batches)
        p_pad = sum((b1 - b0 + 63) // 64 * 64 for b0, b1 in W.batches)
        S = 2
        # algorithmic bytes of the dominant kernel (emit pass): one int64 score per (pod,node) + each table once per launch
        nd_cols, pd_cols = 15 + 2 * S, 8 + S
        emit_bytes = P * N * 8 + nb * N * nd_cols * 8 + P * (pd_cols + 4) * 8
        path_bytes = P * N * (8 + 1 / 8) + nb * 2 * N * nd_cols * 8 + 2 * P * pd_cols * 8 + P * 10 * 8
        peak, peak_src = measured_peak_gbs()
        km = rec["staged"]
        achieved = emit_bytes / (km["emit_ms"] * 1e-3) / 1e9  # emit_ms sums the step's emit launches (one per pod batch)
        traffic, traffic_src = (None, "only quoted for the configuration that was profiled (C2, one GPU)")
        if cfg_id == 2 and world == 1 and nb == 1 and P == base["n_pending"]:
            traffic, traffic_src = emit_traffic_record()
        ms = rec["ms"]
        line = {
            "metric": METRIC, "value": rec["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "int64",
            "data": "synthetic",
            "config": {"workload": workload_name(cfg_id, W.n_total, P, world, scaling), "nodes_per_gpu": N,
                       "l2": "outputs (8 B per eval, %.1f GB per step per GPU) exceed the 126 MB L2; node and pod tables "
                             "(%.1f MB) and the Filter level tables are L2-resident by design" % (P * N * 8 / 1e9, (N * nd_cols * 8 + P * pd_cols * 8) / 1e6),
                       "parity": "all nodes evaluated for every pod (percentageOfNodesToScore=100)",
                       "pod_batches": nb, "merge": W.merge_name, "loop": rec["loop"]},
            "kernel_ms": dict({"filter_pass": km["reduce_ms"], "emit_pass": km["emit_ms"],
                               "note": "per step (all %d batches), measured one batch at a time with every stage bracketed by events; "
                                       "ms_per_step / value: %s" % (nb, rec["loop"])},
                              **({"merge_after_filter_median_per_batch": km["merge1_ms"], "merge_after_emit_median_per_batch": km["merge2_ms"]}
                                 if world > 1 else {}),
                              filter="table look-ups (level tables, filter_rows)" if rec["filter_lookup"] else "compare kernels"),
            "scheduling_cycles_per_sec": args.steps * P / (ms * 1e-3),
            "scheduling_cycles_note": "batched: every pod against the same snapshot, no commit between pods; "
                                      "'sequential' below commits each placement before the next pod",
            "sequential": seq,
            "roofline": {"bound": "hbm", "kernel": "emit_narrow (int32 emit pass)" if rec["emit_int32"] else "eval_kernel<EMIT> (int64 emit pass)",
                         "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "traffic_source": traffic_src, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": emit_bytes // nb, "launches_per_step": nb,
                         "path_frac": (path_bytes / ((ms / args.steps) * 1e-3) / 1e9) / peak,
                         "path_frac_note": "SURVEY §8(d) B_alg of the whole step / ms_per_step / peak",
                         "peak_note": "the peak is the copy figure (read + write); a pure write stream reaches 7.5 TB/s on this GPU and the emit "
                                      "pass's own store pattern without its arithmetic 7.16 TB/s (profiles/r02_write_peak.json, "
                                      "tools/write_pattern.cu, DESIGN.md §14)"},
            "e2e": ({"value": e2e["pipe"]["value"], "unit": UNIT, "h2d_bytes_per_step": pd_cols * P * 8, "d2h_bytes_per_step": 2 * p_pad * 8,
                     "steps": e2e["pipe"]["steps"],
                     "api": "fgs_submit + fgs_collect, %d batches in flight: pinned host pod tables in (fgs_packer_pin: DMA straight from the "
                            "table), per-pod selected node / count / error out; every step copies H2D and reads back D2H" % W.depth
                            + ("" if world == 1 else "; cross-shard merges inside the calls, on their own stream"),
                     "single_call": {"value": e2e["blocking"]["value"], "steps": e2e["blocking"]["steps"],
                                     "api": "fgs_evaluate, one blocking call per batch" if world == 1 else
                                            "fgs_stage_pods + fgs_pass_reduce + merge + fgs_pass_emit + merge + fgs_fetch_results per batch"}}
                    if "pipe" in e2e else
                    {"value": e2e["blocking"]["value"], "unit": UNIT, "h2d_bytes_per_step": pd_cols * P * 8, "d2h_bytes_per_step": 2 * p_pad * 8,
                     "steps": e2e["blocking"]["steps"],
                     "api": "fgs_stage_pods + fgs_pass_reduce + all-reduce + fgs_pass_emit + all-reduce + fgs_fetch_results per batch"}),
            "gpu_launches": rec["launches"],
            "clocks": clocks,
            "parity": parity,
        }
        line.update(extra)
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_port_baseline(fgs, W.synth, W.plugins)
        print(json.dumps(line))
    W.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def run_config3(args, reference):
    """--config 3: Coscheduling's PreFilter for every pending pod of 2 000 PodGroups over 10 000 nodes
    (pkg/coscheduling/core/core.go:149-196, CheckClusterResource :322-342) — PodGroup checks per second"""
    import numpy as np
    import flex_gpu_scheduler_b200 as fgs
    from flex_gpu_scheduler_b200 import _lib as L2
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0  # replicas only: the tables are a few MB, there is nothing to shard
    s = fgs.config_synth(3)
    G, P, N = int(s.podgroups.contents.n_groups), s.n_pending, s.n_nodes
    metric, unit = "podgroup_prefilter_checks_per_sec", "PG-checks/s"
    wl = f"C3: {N} nodes, {G} PodGroups (minMember 4..64), {P} pending pods: PodGroupManager.PreFilter + Permit for every pending pod"
    from oracle import oracle
    cores = os.cpu_count() or 1

    def cpu_sample(budget_s):
        """the oracle's PreFilter on the first pending member of as many groups as the budget allows"""
        pg = np.ctypeslib.as_array(s.pending.contents.pg, shape=(P,))
        first = {}
        for i, g in enumerate(pg):
            if g >= 0 and int(g) not in first:
                first[int(g)] = i
        lib = oracle.load()
        code = np.zeros(1, np.uint8)
        t0 = time.perf_counter()
        n = 0
        for g in sorted(first):
            p = first[g]
            lib.fgs_oracle_gang_prefilter(s.snapshot, s.pending, s.podgroups, p, p + 1, code.ctypes.data_as(L2.p_u8))
            n += 1
            if time.perf_counter() - t0 > budget_s:
                break
        dt = time.perf_counter() - t0
        return n / dt, n, dt

    if reference:
        v, n, dt = cpu_sample(20.0)
        print(json.dumps({"impl": "reference", "metric": metric, "value": v, "unit": unit, "n_gpus": args.gpus, "steps": 1, "warmup": 0,
                          "ms_per_step": 1e3 * dt, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int64",
                          "data": "synthetic", "config": {"workload": wl},
                          "cpu_baseline": {"value": v, "unit": unit, "cores": cores, "threads": 1, "kind": "port",
                                           "sample": f"PreFilter of the first pending member of {n} groups, {dt:.1f} s"},
                          "e2e": {"value": v, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}))
        return 0
    import torch
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    eng = fgs.Engine(device=local)
    eng.upload_nodes(fgs.Packer(2).nodes(s.snapshot))
    for _ in range(args.warmup):
        eng.gang_check(s.snapshot, s.pending, s.podgroups)
    torch.cuda.synchronize()
    l0 = eng.timings()["kernel_launches"]
    t0 = time.perf_counter()
    for _ in range(args.steps):
        pre, permit, assigned = eng.gang_check(s.snapshot, s.pending, s.podgroups)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    launches = eng.timings()["kernel_launches"] - l0
    # one check per (pending pod with a PodGroup): that is what the reference runs, one PreFilter per scheduling cycle
    n_checks = int((np.ctypeslib.as_array(s.pending.contents.pg, shape=(P,)) >= 0).sum())
    v = args.steps * n_checks / dt
    cv, cn, cdt = cpu_sample(10.0)
    print(json.dumps({"metric": metric, "value": v, "unit": unit, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
                      "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int64",
                      "data": "synthetic",
                      "config": {"workload": wl, "checks_per_step": n_checks, "replicas": "single GPU (the tables are a few MB: nothing shards)"},
                      "e2e": {"value": v, "unit": unit, "h2d_bytes_per_step": None, "d2h_bytes_per_step": 2 * P + 4 * G,
                              "api": "fgs_gang_check: host object tables in (packed and copied inside the call), per-pod PreFilter / Permit codes out; "
                                     "value IS the end-to-end figure, timed on the host around the blocking call"},
                      "gpu_launches": int(launches),
                      "cpu_baseline": {"value": cv, "unit": unit, "cores": cores, "threads": 1, "kind": "port",
                                       "sample": f"oracle PreFilter (CheckClusterResource walk included) of the first pending member of {cn} groups, {cdt:.1f} s"},
                      "roofline": {"bound": "latency", "note": "a few MB of tables, L2-resident; per call: host packing of the pods on nodes + prefix scan + one CTA per PodGroup"},
                      "codes_seen": sorted(int(x) for x in np.unique(pre))}))
    eng.close()
    return 0


def run_config4(args, reference):
    """--config 4: CapacityScheduling in front of the FlexGPU / Fit path on 20 000 nodes, 500 namespaces, 100 000 pods. One
    step = what the reference's cycles do for the whole pending set: the ElasticQuota PreFilter of every pod
    (pkg/capacityscheduling/capacity_scheduling.go:201-275), then Filter + Score + selectHost of the pods it admits."""
    import ctypes as C
    import numpy as np
    import flex_gpu_scheduler_b200 as fgs
    from flex_gpu_scheduler_b200 import _lib as L2
    from oracle import oracle
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0  # the quota tables do not shard; the evaluation half is what configs 2 / 5 measure sharded
    s = fgs.config_synth(4)
    cfg = fgs.config_plugins(4)
    N, P = s.n_nodes, s.n_pending
    cores = os.cpu_count() or 1
    wl = (f"C4: {N} nodes x 8 GPUs, 500 namespaces with ElasticQuotas, {P} pending pods: CapacityScheduling.PreFilter for every pod, "
          "then FlexGPU+NodeResourcesFit filter and FlexGPU+LeastAllocated scores for the admitted ones, single GPU")
    ref_code = oracle.quota_prefilter(s.pending, s.quotas)
    admitted = np.nonzero(ref_code == 0)[0]

    def cpu_leg(budget_s):
        t0 = time.perf_counter()
        oracle.quota_prefilter(s.pending, s.quotas)
        t_q = time.perf_counter() - t0
        n, t_e = 0, 0.0
        blk = 64
        t0 = time.perf_counter()
        while t_e < budget_s and n < len(admitted):
            p = int(admitted[n])  # contiguous ranges of the pending set starting at admitted pods: same per-pod cost
            oracle.evaluate(s.snapshot, s.pending, cfg, p, min(P, p + blk), workers=cores, planes=())
            n += blk
            t_e = time.perf_counter() - t0
        evals = min(n, P) * N
        # scale the quota walk to the sample: it is per pod too
        return evals / (t_e + t_q * min(n, P) / P), n, t_q, t_e

    if reference:
        v, n, t_q, t_e = cpu_leg(30.0)
        print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": 1, "warmup": 0,
                          "ms_per_step": 1e3 * (t_e + t_q), "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int64",
                          "data": "synthetic", "config": {"workload": wl},
                          "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "threads": cores, "kind": "port",
                                           "sample": f"quota PreFilter of all {P} pods ({t_q:.2f} s, scaled to the sample) + evaluation of {n} pods x all {N} nodes ({t_e:.1f} s)"},
                          "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}))
        return 0
    import torch
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    eng = fgs.Engine(device=local, n_scalars=2, total_nodes=N)
    eng.set_plugins(cfg)
    eng.upload_nodes(fgs.Packer(2).nodes(s.snapshot))
    pk_all = fgs.Packer(2)
    t_all = pk_all.pods(s.pending)

    def gather_table(idx):
        """the pod table of the admitted pods: the columns of the full table gathered (what the Go shim packs after PreFilter)"""
        src = t_all.contents
        keep = []
        t = L2.PodTable()
        t.n_pods, t.n_scalars = len(idx), 2

        def col(ptr, ctype, width=1):
            a = np.ctypeslib.as_array(ptr, shape=(width, P) if width > 1 else (P,))
            g = np.ascontiguousarray(a[..., idx])
            keep.append(g)
            return g.ctypes.data_as(C.POINTER(ctype))
        for name, ct in (("flex_gpu_limit", C.c_int64), ("flex_mem_limit", C.c_int64), ("req_mcpu", C.c_int64), ("req_mem", C.c_int64),
                         ("req_eph", C.c_int64), ("nz_mcpu", C.c_int64), ("nz_mem", C.c_int64), ("acct_nz_mcpu", C.c_int64),
                         ("acct_nz_mem", C.c_int64)):
            setattr(t, name, col(getattr(src, name), ct))
        t.req_scalar = col(src.req_scalar, C.c_int64, 2)
        for name, ct in (("flags", C.c_uint32), ("req_scalar_present", C.c_uint32), ("ns", C.c_int32), ("pg", C.c_int32), ("priority", C.c_int32),
                         ("index_annot", C.c_int32)):
            setattr(t, name, col(getattr(src, name), ct))
        p = C.pointer(t)
        p._keep = keep
        return p

    def step():
        code = eng.quota_check(t_all, s.quotas)  # every pod's PreFilter
        idx = np.nonzero(code == 0)[0]
        res = eng.evaluate(gather_table(idx))    # the admitted pods go on to Filter / Score / selectHost
        return code, idx, res

    for _ in range(args.warmup):
        code, idx, res = step()
    torch.cuda.synchronize()
    l0 = eng.timings()["kernel_launches"]
    t0 = time.perf_counter()
    tq = 0.0
    for _ in range(args.steps):
        t1 = time.perf_counter()
        code = eng.quota_check(t_all, s.quotas)
        tq += time.perf_counter() - t1
        idx = np.nonzero(code == 0)[0]
        res = eng.evaluate(gather_table(idx))
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    launches = eng.timings()["kernel_launches"] - l0
    tm = eng.timings()
    assert np.array_equal(code, ref_code), "quota verdicts differ from the oracle"
    n_adm = len(idx)
    v = args.steps * n_adm * N / dt
    cv, cn, ct_q, ct_e = cpu_leg(10.0)
    peak, peak_src = measured_peak_gbs()
    emit_bytes = n_adm * N * 8
    print(json.dumps({"metric": METRIC, "value": v, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
                      "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int64",
                      "data": "synthetic",
                      "config": {"workload": wl, "admitted_pods": int(n_adm), "rejected_over_max": int((code == 1).sum()),
                                 "rejected_over_min": int((code == 2).sum()), "evals_per_step": int(n_adm * N)},
                      "kernel_ms": {"quota_check_call_ms": 1e3 * tq / args.steps, "filter_pass": tm["reduce_ms"], "emit_pass": tm["emit_ms"],
                                    "h2d": tm["h2d_ms"], "note": "quota_check_call_ms is the whole blocking call (host tables in, verdicts out); "
                                                                 "the passes are those of the last evaluation"},
                      "roofline": {"bound": "hbm", "kernel": "emit_narrow (int32 emit pass)", "achieved": emit_bytes / (tm["emit_ms"] * 1e-3) / 1e9,
                                   "peak": peak, "unit": "GB/s", "frac": emit_bytes / (tm["emit_ms"] * 1e-3) / 1e9 / peak, "traffic": None,
                                   "peak_source": peak_src, "algorithmic_bytes_per_launch": emit_bytes},
                      "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": int(10 * n_adm * 8 + P * 8 * 4), "d2h_bytes_per_step": int(16 * n_adm + P),
                              "api": "fgs_quota_check (host tables in, verdicts out) + host gather of the admitted pods + fgs_evaluate (host pod table in, "
                                     "per-pod results out), blocking calls: value IS the end-to-end figure"},
                      "gpu_launches": int(launches),
                      "cpu_baseline": {"value": cv, "unit": UNIT, "cores": cores, "threads": cores, "kind": "port",
                                       "sample": f"quota PreFilter of all {P} pods ({ct_q:.2f} s, scaled) + evaluation of {cn} pods x all {N} nodes ({ct_e:.1f} s)"},
                      "parity": {"quota_verdicts_equal_oracle": True}}))
    eng.close()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", type=int, default=0, help="1..5 (default: 2 on one GPU, 5 node-sharded on several)")
    ap.add_argument("--pods", type=int, default=0, help="override the number of pending pods")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sequential", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle spot check of the pipeline's results")
    ap.add_argument("--no-c5", action="store_true", help="N = 1: skip the config-5 single-GPU record")
    ap.add_argument("--no-weak", action="store_true", help="N > 1: skip the config-2 weak-scaling record")
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

"""cross_fuzz.py — fuzz the two CPU restatements against each other. TEST INFRASTRUCTURE ONLY.

oracle/ref_walk.cpp (C++, the checker of the GPU parity tests and the timed CPU baseline) and oracle/literal_walk.py
(Python, written from the Go independently) evaluate the same random small snapshots; every output they share must
agree: Filter status per (pod,node) — code and failing plugin —, raw and normalised scores per plugin on feasible
nodes, weighted totals, feasible counts, cycle-error bits, selectHost's maximum and its tie set, FlexGPU.Reserve's
device index. The reference holds no test for these functions, so this does not PIN either of them; it shows that two
separate readings of the Go give the same answers.

    python -m oracle.cross_fuzz --n 100000 [--seed 1]
"""
import argparse
import os
import random
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import builders as B  # noqa: E402
import flex_gpu_scheduler_b200 as fgs  # noqa: E402
from oracle import literal_walk as LW  # noqa: E402
from oracle import oracle as O  # noqa: E402

MI = 1 << 20


def rand_requests(rng, S, allow_missing=True):
    req = {}
    if not (allow_missing and rng.random() < 0.2):
        req["cpu"] = rng.choice([0, 1, 50, 100, 250, 999, 1000, 1001, 2500])
    if not (allow_missing and rng.random() < 0.2):
        req["memory"] = rng.choice([0, 1, 64, 200, 256, 1024, 4096]) * MI
    if rng.random() < 0.15:
        req["ephemeral-storage"] = rng.choice([0, 1, 10, 100]) * MI
    for s in range(2, S):
        if rng.random() < 0.2:
            req[("scalar", s)] = rng.choice([0, 1, 2, 5])
    return req


def rand_pod(rng, S, on_node=False, n_dev=4, mem_each=16384):
    kind = rng.choices([0, 1, 2, 3], weights=[2, 3, 5, 1])[0]
    ctrs = []
    n_ctr = rng.choice([1, 1, 1, 2, 3])
    for c in range(n_ctr):
        req = rand_requests(rng, S)
        gl = ml = None
        if c == 0 or rng.random() < 0.2:
            if kind in (1, 3):
                gl = rng.choice([0, 1, 1, 1, 2])
                if rng.random() < 0.8:
                    req[("scalar", 0)] = gl
            if kind in (2, 3):
                ml = rng.choice([0, 1, mem_each // 4, mem_each // 2, mem_each, mem_each + 1, 1000, 3000])
                if rng.random() < 0.8:
                    req[("scalar", 1)] = ml
        ctrs.append(B.ctr(req, gpu_limit=gl, mem_limit=ml))
    inits = [B.ctr(rand_requests(rng, S, allow_missing=rng.random() < 0.5)) for _ in range(rng.choice([0, 0, 0, 1, 2]))]
    ovh = None
    if rng.random() < 0.2:
        ovh = {}
        if rng.random() < 0.8:
            ovh["cpu"] = rng.choice([0, 1, 125, 500, 1000, 1001, 1999, 2000])
        if rng.random() < 0.8:
            ovh["memory"] = rng.choice([0, 1, 64 * MI])
        if rng.random() < 0.1:
            ovh["ephemeral-storage"] = MI
    index = None
    if on_node and kind != 0:
        r = rng.random()
        if r < 0.7:
            index = rng.randrange(0, max(n_dev, 1))
        elif r < 0.8:
            index = rng.choice([-1, n_dev, n_dev + 3, 100])  # NewGPUNode indexes out of range
    elif rng.random() < 0.05:
        index = rng.randrange(0, 4)
    return B.pod(containers=ctrs, init_containers=inits, overhead=ovh, node_name=on_node, index=index)


def rand_node(rng, S):
    n_dev = rng.choice([1, 2, 4, 4, 8, 8, 11])
    mem_each = rng.choice([8192, 16384, 16385, 24576])
    alloc = {"cpu": rng.choice([0, 4000, 8000, 32000, 32000, 64000]), "memory": rng.choice([0, 8192, 65536, 262144, 262144]) * MI,
             "ephemeral-storage": rng.choice([0, 100, 1000, 1000]) * MI, "pods": rng.choice([0, 1, 3, 10, 110, 110])}
    r = rng.random()
    if r < 0.75:
        alloc[("scalar", 0)] = n_dev
        alloc[("scalar", 1)] = n_dev * mem_each + rng.choice([0, 0, 1, 3])
    elif r < 0.82:  # quirks: a zero or negative device count, one key missing
        alloc[("scalar", 0)] = rng.choice([0, -1])
        alloc[("scalar", 1)] = mem_each
    elif r < 0.88:
        alloc[("scalar", 0)] = n_dev
    elif r < 0.92:
        alloc[("scalar", 1)] = mem_each
    for s in range(2, S):
        if rng.random() < 0.5:
            alloc[("scalar", s)] = rng.choice([0, 1, 4, 16])
    pods = [rand_pod(rng, S, on_node=True, n_dev=n_dev, mem_each=mem_each) for _ in range(rng.choice([0, 0, 1, 2, 3, 5]))]
    nd = B.node(alloc, pods, valid=rng.random() > 0.04)
    if rng.random() < 0.15:  # aggregates that are not the sum of the listed pods
        nd["req_bias"] = [rng.choice([0, 100, -100]), rng.choice([0, MI]), 0] + [0] * S
        nd["nz_bias"] = [rng.choice([0, 100]), rng.choice([0, MI])]
    return nd


def rand_cfg(rng, S):
    filt = rng.choice([("NodeResourcesFit", "FlexGPU"), ("FlexGPU", "NodeResourcesFit"), ("FlexGPU",), ("NodeResourcesFit",)])
    names = ["FlexGPU", "NodeResourcesFit", "NodeResourcesAllocatable"]
    k = rng.choice([1, 2, 2, 3])
    score = tuple((nm, rng.choice([0, 1, 2, 5])) for nm in rng.sample(names, k))
    res_pool = [("cpu", rng.randint(1, 9)), ("memory", rng.randint(1, 9))]
    if rng.random() < 0.3:
        res_pool.append(("ephemeral-storage", rng.randint(1, 5)))
    for s in range(S):
        if rng.random() < 0.2:
            res_pool.append((("scalar", s), rng.randint(1, 5)))
    fit_res = tuple(rng.sample(res_pool, rng.randint(1, len(res_pool))))
    alloc_res = tuple((nm, rng.choice([1, 3, 1 << 20])) for nm in rng.sample(["cpu", "memory", "ephemeral-storage", ("scalar", 1), "pods"],
                                                                             rng.randint(1, 3)))
    kw = dict(filter_plugins=filt, score_plugins=score, fit_strategy=rng.choice(["LeastAllocated", "MostAllocated"]), fit_resources=fit_res,
              alloc_mode=rng.choice(["Least", "Most"]), alloc_resources=alloc_res)
    if S > 2 and rng.random() < 0.3:
        kw["fit_ignored_scalar_mask"] = 1 << rng.randrange(2, S)
    return fgs.plugin_cfg(**kw)


def compare(snap, pend, cfg, S, n_nodes, n_pods):
    """→ list of mismatch descriptions (empty = the two restatements agree)"""
    ref = O.evaluate(snap.ptr, pend.ptr, cfg, workers=1, per_plugin=True)
    lit = LW.evaluate(snap.ptr, pend.ptr, cfg)
    bad = []
    sm = int(cfg.score_mask)
    for p in range(n_pods):
        r = lit[p]
        feas_ref = ref.status[p] == 0
        feas_lit = np.zeros(n_nodes, bool)
        feas_lit[r.feasible] = True
        if not np.array_equal(feas_ref, feas_lit):
            bad.append(f"pod {p}: feasible sets differ {feas_ref.astype(int)} vs {feas_lit.astype(int)}")
            continue
        for n in range(n_nodes):
            a, b = int(ref.status[p, n]), int(r.status[n])
            if b == 0xFF:  # a row that is not a node: only "not feasible" is defined
                continue
            if b == 0xF0:
                if (a & 0xF0) != 0xF0:
                    bad.append(f"pod {p} node {n}: literal panics in Filter, ref_walk status {a:#x}")
            elif a != b:
                bad.append(f"pod {p} node {n}: status {a:#x} vs {b:#x}")
        if int(ref.feasible_count[p]) != len(r.feasible):
            bad.append(f"pod {p}: count {ref.feasible_count[p]} vs {len(r.feasible)}")
        if int(ref.pod_err[p]) != r.pod_err:
            bad.append(f"pod {p}: pod_err {ref.pod_err[p]} vs {r.pod_err}")
        if r.pod_err or not r.feasible:
            continue
        f = r.feasible
        for pid, raw, norm in ((LW.PL_FLEXGPU, ref.raw_flex, ref.norm_flex), (LW.PL_FIT, ref.raw_fit, ref.norm_fit),
                               (LW.PL_ALLOCATABLE, ref.raw_alloc, ref.norm_alloc)):
            if not (sm & pid):
                continue
            if not np.array_equal(raw[p, f], r.raw[pid][f]):
                bad.append(f"pod {p}: raw scores of plugin {pid} differ {raw[p, f]} vs {r.raw[pid][f]}")
            if not np.array_equal(norm[p, f], r.norm[pid][f]):
                bad.append(f"pod {p}: normalised scores of plugin {pid} differ {norm[p, f]} vs {r.norm[pid][f]}")
        if not np.array_equal(ref.total[p, f], r.total[f]):
            bad.append(f"pod {p}: totals differ {ref.total[p, f]} vs {r.total[f]}")
        if int(ref.max_total[p]) != r.max_total or int(ref.tie_count[p]) != len(r.tie_set) or int(ref.first_max[p]) != r.tie_set[0]:
            bad.append(f"pod {p}: selectHost {ref.max_total[p]}/{ref.tie_count[p]}/{ref.first_max[p]} vs {r.max_total}/{len(r.tie_set)}/{r.tie_set[0]}")
    # FlexGPU.Reserve's device choice on a few (pod,node) pairs
    for p in range(min(n_pods, 3)):
        for n in range(min(n_nodes, 3)):
            if not int(snap.struct.node_valid[n]):
                continue
            try:
                b = LW.reserve_index(snap.ptr, pend.ptr, p, n)
            except NotImplementedError:
                continue
            a = O.reserve_index(snap.ptr, pend.ptr, p, n)
            if a != b:
                bad.append(f"Reserve pod {p} node {n}: {a} vs {b}")
    return bad


def one_case(rng):
    S = rng.choice([2, 2, 2, 3, 4])
    n_nodes, n_pods = rng.randint(1, 6), rng.randint(1, 5)
    nodes = [rand_node(rng, S) for _ in range(n_nodes)]
    pods = [rand_pod(rng, S) for _ in range(n_pods)]
    snap, pend = B.build_snapshot(nodes, S), B.build_podset(pods, S)
    cfg = rand_cfg(rng, S)
    return compare(snap, pend, cfg, S, n_nodes, n_pods), (nodes, pods, cfg)


def run(n, seed=1, stop_at=5, quiet=False):
    rng = random.Random(seed)
    O.load()
    t0 = time.time()
    n_bad = 0
    evals = 0
    for i in range(n):
        bad, (nodes, pods, cfg) = one_case(rng)
        evals += len(nodes) * len(pods)
        if bad:
            n_bad += 1
            if not quiet:
                print(f"case {i} (seed {seed}): {len(bad)} mismatches")
                for b in bad[:6]:
                    print("   ", b)
            if n_bad >= stop_at:
                break
    dt = time.time() - t0
    if not quiet:
        print(f"{i + 1} snapshots, {evals} (pod,node) evaluations, {n_bad} snapshots with a mismatch, {dt:.1f} s, seed {seed}")
    return n_bad, i + 1, evals


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=2000)
    ap.add_argument("--seed", type=int, default=1)
    a = ap.parse_args()
    sys.exit(1 if run(a.n, a.seed)[0] else 0)

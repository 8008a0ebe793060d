"""Reference side of the sequential-cycle tests: "the oracle follows the engine's tie choice" (SURVEY.md §8 f1).

For every pod of the queue, in order, the oracle evaluates the pod against a snapshot rebuilt from the CURRENT
Python description of the cluster; the engine's choice must lie in the oracle's max-score tie set; then the
reference's assume + Reserve are applied to the description (the pod joins the node's pod list with the index
annotation FlexGPU.Reserve sets — scheduler.go:496-521, flex_gpu.go:178-223) and the next pod sees it.
NodeInfo aggregates are re-accumulated by builders.build_snapshot the way NodeInfo.AddPod does."""
import copy
import ctypes as C

import numpy as np

import builders as B
import flex_gpu_scheduler_b200._lib as L

RESERVE_FAIL, RESERVE_NONE, RESERVE_CONFLICT, RESERVE_PANIC, RESERVE_SKIPPED = -1, -2, -3, -4, -5


def _arr(ptr, n, dtype=None):
    if n == 0:
        return np.zeros(0, dtype or np.int64)
    return np.ctypeslib.as_array(ptr, shape=(n,))


def podset_to_pods(ps):
    """flat fgs_podset → list of builders.pod dicts"""
    n, S = int(ps.n_pods), int(ps.n_scalars)
    W = 3 + S
    if n == 0:
        return []
    ctr_off = _arr(ps.ctr_off, n + 1)
    C_ = int(ctr_off[-1])
    c_req = _arr(ps.c_req, max(C_, 1) * W).reshape(-1, W)
    c_pres = _arr(ps.c_req_present, max(C_, 1))
    c_lim = _arr(ps.c_lim_flex, max(C_, 1) * 2).reshape(-1, 2)
    c_lp = _arr(ps.c_lim_present, max(C_, 1))
    n_init = _arr(ps.n_init, n) if ps.n_init else np.zeros(n, np.int32)
    ovp = _arr(ps.overhead_present, n) if ps.overhead_present else np.zeros(n, np.uint32)
    ov = _arr(ps.overhead, n * W).reshape(-1, W) if ps.overhead else np.zeros((n, W), np.int64)
    ns, pg, prio = _arr(ps.ns, n), _arr(ps.pg, n), _arr(ps.priority, n)
    named, annot = _arr(ps.has_node_name, n), _arr(ps.index_annot, n)
    keys = ["cpu", "memory", "ephemeral-storage"] + [("scalar", s) for s in range(S)]

    def ctr(k):
        req = {keys[r]: int(c_req[k, r]) for r in range(W) if (int(c_pres[k]) >> r) & 1}
        lp = int(c_lp[k])
        return B.ctr(req, gpu_limit=int(c_lim[k, 0]) if lp & 1 else None, mem_limit=int(c_lim[k, 1]) if lp & 2 else None)

    pods = []
    for i in range(n):
        a, b = int(ctr_off[i]), int(ctr_off[i + 1])
        ni = int(n_init[i])
        regular = [ctr(k) for k in range(a, b - ni)]
        init = [ctr(k) for k in range(b - ni, b)]
        overhead = None
        if int(ovp[i]) >> 31:
            overhead = {keys[r]: int(ov[i, r]) for r in range(W) if (int(ovp[i]) >> r) & 1}
        idx = int(annot[i])
        pods.append(B.pod(regular, init, overhead, ns=int(ns[i]), pg=int(pg[i]), priority=int(prio[i]), node_name=bool(named[i]),
                          index=None if idx == L.FGS_ANNOT_NONE else idx))
    return pods


def snapshot_to_nodes(sn):
    """flat fgs_snapshot → list of builders.node dicts (pods attached)"""
    N, S = int(sn.n_nodes), int(sn.n_scalars)
    AW = 4 + S
    alloc = _arr(sn.alloc, N * AW).reshape(N, AW)
    present = _arr(sn.alloc_scalar_present, N)
    valid = _arr(sn.node_valid, N)
    pod_off = _arr(sn.pod_off, N + 1)
    pods = podset_to_pods(sn.pods)
    req = _arr(sn.requested, N * (3 + S)).reshape(N, 3 + S)
    nz = _arr(sn.nonzero, N * 2).reshape(N, 2)
    nodes = []
    for n in range(N):
        a = {"cpu": int(alloc[n, 0]), "memory": int(alloc[n, 1]), "ephemeral-storage": int(alloc[n, 2]), "pods": int(alloc[n, 3])}
        for s in range(S):
            if (int(present[n]) >> s) & 1:
                a[("scalar", s)] = int(alloc[n, 4 + s])
        nd = B.node(a, pods[int(pod_off[n]):int(pod_off[n + 1])], valid=bool(valid[n]))
        # the generator's quirk nodes carry aggregates that differ from the sum over their pods: keep the difference
        rs, cs, ms = [0] * (3 + S), 0, 0
        for p in nd["pods"]:
            r, c, m = B._calc_resource(p, S)
            rs = [x + y for x, y in zip(rs, r)]
            cs, ms = cs + c, ms + m
        nd["req_bias"] = [int(req[n, k]) - rs[k] for k in range(3 + S)]
        nd["nz_bias"] = [int(nz[n, 0]) - cs, int(nz[n, 1]) - ms]
        nodes.append(nd)
    return nodes


class QuotaState:
    """Mutable copy of an fgs_quotas table (ElasticQuotaInfos) for the stepped reference."""

    def __init__(self, q):
        NS, S = int(q.n_ns), int(q.n_scalars)
        W = 2 + S
        self.NS, self.S = NS, S
        self.has_eq = np.array(_arr(q.has_eq, NS), np.uint8).copy()
        self.min = np.array(_arr(q.min, NS * W), np.int64).reshape(NS, W).copy()
        self.max = np.array(_arr(q.max, NS * W), np.int64).reshape(NS, W).copy()
        self.used = np.array(_arr(q.used, NS * W), np.int64).reshape(NS, W).copy()
        self.present = np.array(_arr(q.used_scalar_present, NS), np.uint32).copy()

    def struct(self):
        q = L.Quotas()
        q.n_ns, q.n_scalars = self.NS, self.S
        q.has_eq = self.has_eq.ctypes.data_as(L.p_u8)
        q.min = self.min.ctypes.data_as(L.p_i64)
        q.max = self.max.ctypes.data_as(L.p_i64)
        q.used = self.used.ctypes.data_as(L.p_i64)
        q.used_scalar_present = self.present.ctypes.data_as(L.p_u32)
        return q


def follow(oracle, nodes, queue, cfg, res, reserve_flex=True, S=2, check_scores=True, quotas=None):
    """Step the reference through `queue` following the engine's choices in `res` (api.SeqResults).
    Mutates and returns `nodes` (the cluster after every commit). Asserts parity at every cycle."""
    stats = dict(committed=0, unschedulable=0, reserve_fail=0, errors=0, quota_rejected=0)
    for p, pd in enumerate(queue):
        snap = B.build_snapshot(nodes, S)
        pend = B.build_podset([pd], S)
        if quotas is not None:
            # CapacityScheduling.PreFilter against the ElasticQuotaInfos as the previous commits left them
            qs = quotas.struct()
            qcode = int(oracle.quota_prefilter(pend.ptr, C.pointer(qs))[0])
            assert int(res.quota_code[p]) == qcode, (p, int(res.quota_code[p]), qcode)
            if qcode:
                assert int(res.feasible_count[p]) == 0 and int(res.selected_node[p]) == -1 and not res.committed[p], p
                stats["quota_rejected"] += 1
                continue
        ref = oracle.evaluate(snap.ptr, pend.ptr, cfg, 0, 1)
        feas = ref.status[0] == 0
        F = int(feas.sum())
        assert int(res.feasible_count[p]) == F, (p, int(res.feasible_count[p]), F)
        assert int(res.pod_err[p]) & 3 == int(ref.pod_err[0]) or int(res.device_index[p]) == RESERVE_PANIC, (p, res.pod_err[p], ref.pod_err[0])
        sel = int(res.selected_node[p])
        if F == 0:
            assert sel == -1 and not res.committed[p] and int(res.device_index[p]) == RESERVE_SKIPPED, p
            stats["unschedulable"] += 1
            continue
        assert 0 <= sel < len(nodes) and feas[sel], (p, sel)
        if int(ref.pod_err[0]):
            assert not res.committed[p] and int(res.device_index[p]) == RESERVE_SKIPPED, p
            stats["errors"] += 1
            continue
        if F >= 2:
            assert int(ref.total[0][sel]) == int(ref.max_total[0]), (p, sel, int(ref.total[0][sel]), int(ref.max_total[0]))
            if check_scores:
                assert int(res.selected_score[p]) == int(ref.max_total[0]), p
        want = RESERVE_SKIPPED
        ok = True
        if reserve_flex:
            want = oracle.reserve_index(snap.ptr, pend.ptr, 0, sel)
            ok = want >= 0 or want == RESERVE_NONE
        assert int(res.device_index[p]) == want, (p, int(res.device_index[p]), want)
        assert bool(res.committed[p]) == ok, p
        if not ok:
            stats["reserve_fail"] += 1
            continue
        placed = copy.deepcopy(pd)
        placed["node_name"] = True
        if want >= 0:
            placed["index"] = want  # p.Annotations[GPUIndexAnnotationKey] = strconv.Itoa(indexes[0])
        nodes[sel]["pods"].append(placed)
        stats["committed"] += 1
        if quotas is not None:  # CapacityScheduling.Reserve → addPodIfNotPresent → reserveResource
            ns = pd["ns"]
            if 0 <= ns < quotas.NS and quotas.has_eq[ns]:
                quotas.present[ns] = oracle.quota_reserve(pend.ptr, 0, quotas.used[ns], int(quotas.present[ns]))
    return nodes, stats

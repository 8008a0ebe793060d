"""Fixture builders: Python dict descriptions of nodes / pods / groups / quotas → the flat object
snapshot of include/fgs_types.h. They play the role of the reference's test helpers
(st.MakeNode().Capacity, st.MakePod().Req, testutil.NewFakeSharedLister — test/util/fake.go:38-98):
NodeInfo aggregates are accumulated the way NodeInfo.AddPod does (framework/types.go:583-609,715-745).

Resource keys: "cpu" (milli-cores), "memory", "ephemeral-storage", "pods", ("scalar", s).
"""
import ctypes as C

import numpy as np

import flex_gpu_scheduler_b200._lib as L

ANNOT_NONE = L.FGS_ANNOT_NONE
DEFAULT_CPU, DEFAULT_MEM = 100, 200 * 1024 * 1024


def _slot(key):
    if key == "cpu":
        return 0
    if key == "memory":
        return 1
    if key == "ephemeral-storage":
        return 2
    if isinstance(key, tuple) and key[0] == "scalar":
        return 3 + key[1]
    raise KeyError(key)


def pod(containers=None, init_containers=None, overhead=None, ns=0, pg=-1, priority=0, node_name=False, index=None, uid=None, start=0):
    """containers: list of {"requests": {...}, "limits": {"gpu": v, "mem": v}}"""
    return dict(containers=containers or [], init_containers=init_containers or [], overhead=overhead, ns=ns, pg=pg,
                priority=priority, node_name=node_name, index=index, uid=uid, start=start)


def ctr(requests=None, gpu_limit=None, mem_limit=None):
    return dict(requests=requests or {}, gpu_limit=gpu_limit, mem_limit=mem_limit)


class Built:
    """keeps the numpy arrays alive next to the ctypes struct"""

    def __init__(self, struct, keep):
        self.struct = struct
        self.keep = keep

    @property
    def ptr(self):
        return C.pointer(self.struct)


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def build_podset(pods, S=2):
    W = 3 + S
    n = len(pods)
    ns = np.array([p["ns"] for p in pods] or [0], np.int32)
    pg = np.array([p["pg"] for p in pods] or [0], np.int32)
    prio = np.array([p["priority"] for p in pods] or [0], np.int32)
    named = np.array([1 if p["node_name"] else 0 for p in pods] or [0], np.uint8)
    annot = np.array([ANNOT_NONE if p["index"] is None else p["index"] for p in pods] or [0], np.int32)
    start = np.array([p.get("start", 0) for p in pods] or [0], np.int64)
    ctr_off = [0]
    n_init = []
    c_req, c_pres, c_lim, c_lpres = [], [], [], []
    ovh = np.zeros((max(n, 1), W), np.int64)
    ovh_p = np.zeros(max(n, 1), np.uint32)
    for i, p in enumerate(pods):
        for c in list(p["containers"]) + list(p["init_containers"]):
            row = [0] * W
            pres = 0
            for k, v in c["requests"].items():
                s = _slot(k)
                row[s] = v
                pres |= 1 << s
            c_req.append(row)
            c_pres.append(pres)
            lp = 0
            gl = ml = 0
            if c.get("gpu_limit") is not None:
                gl = c["gpu_limit"]
                lp |= 1
            if c.get("mem_limit") is not None:
                ml = c["mem_limit"]
                lp |= 2
            c_lim.append([gl, ml])
            c_lpres.append(lp)
        ctr_off.append(len(c_req))
        n_init.append(len(p["init_containers"]))
        if p["overhead"] is not None:
            ovh_p[i] = 0x80000000
            for k, v in p["overhead"].items():
                s = _slot(k)
                ovh[i, s] = v
                ovh_p[i] |= 1 << s
    ctr_off = np.array(ctr_off, np.int64)
    n_init = np.array(n_init or [0], np.int32)
    c_req = np.array(c_req or [[0] * W], np.int64).reshape(-1, W)
    c_pres = np.array(c_pres or [0], np.uint32)
    c_lim = np.array(c_lim or [[0, 0]], np.int64).reshape(-1, 2)
    c_lpres = np.array(c_lpres or [0], np.uint32)
    ps = L.Podset()
    ps.n_pods = n
    ps.n_scalars = S
    ps.ns, ps.pg, ps.priority = _p(ns, C.c_int32), _p(pg, C.c_int32), _p(prio, C.c_int32)
    ps.has_node_name, ps.index_annot = _p(named, C.c_uint8), _p(annot, C.c_int32)
    ps.ctr_off, ps.n_init = _p(ctr_off, C.c_int64), _p(n_init, C.c_int32)
    ps.overhead_present, ps.overhead = _p(ovh_p, C.c_uint32), _p(ovh, C.c_int64)
    ps.n_ctrs = len(c_pres) if pods and any(p["containers"] or p["init_containers"] for p in pods) else 0
    ps.c_req, ps.c_req_present = _p(c_req, C.c_int64), _p(c_pres, C.c_uint32)
    ps.c_lim_flex, ps.c_lim_present = _p(c_lim, C.c_int64), _p(c_lpres, C.c_uint32)
    ps.start_time = _p(start, C.c_int64)
    return Built(ps, [ns, pg, prio, named, annot, ctr_off, n_init, ovh, ovh_p, c_req, c_pres, c_lim, c_lpres, start])


def _calc_resource(p, S):
    """calculateResource (framework/types.go:715-745) for the fixture's NodeInfo aggregates"""
    W = 3 + S
    res = [0] * W
    n0c = n0m = 0
    for c in p["containers"]:
        for k, v in c["requests"].items():
            res[_slot(k)] += v
        n0c += c["requests"].get("cpu", DEFAULT_CPU)
        n0m += c["requests"].get("memory", DEFAULT_MEM)
    for c in p["init_containers"]:
        for k, v in c["requests"].items():
            res[_slot(k)] = max(res[_slot(k)], v)
        n0c = max(n0c, c["requests"].get("cpu", DEFAULT_CPU))
        n0m = max(n0m, c["requests"].get("memory", DEFAULT_MEM))
    if p["overhead"] is not None:
        for k, v in p["overhead"].items():
            res[_slot(k)] += v
        n0c += p["overhead"].get("cpu", 0)
        n0m += p["overhead"].get("memory", 0)
    return res, n0c, n0m


def node(alloc=None, pods=None, valid=True):
    return dict(alloc=alloc or {}, pods=pods or [], valid=valid)


def build_snapshot(nodes, S=2):
    N = len(nodes)
    AW, RW = 4 + S, 3 + S
    alloc = np.zeros((max(N, 1), AW), np.int64)
    present = np.zeros(max(N, 1), np.uint32)
    req = np.zeros((max(N, 1), RW), np.int64)
    nz = np.zeros((max(N, 1), 2), np.int64)
    valid = np.ones(max(N, 1), np.uint8)
    pod_off = [0]
    all_pods = []
    for i, nd in enumerate(nodes):
        valid[i] = 1 if nd["valid"] else 0
        for k, v in nd["alloc"].items():
            if k == "pods":
                alloc[i, 3] = v
            elif isinstance(k, tuple):
                alloc[i, 4 + k[1]] = v
                present[i] |= 1 << k[1]
            else:
                alloc[i, _slot(k)] = v
        # aggregates that are not the sum of the listed pods (a cache that is ahead of / behind the pod list)
        req[i] += np.array(nd.get("req_bias", [0] * RW), np.int64)
        nz[i] += np.array(nd.get("nz_bias", [0, 0]), np.int64)
        for p in nd["pods"]:
            res, c, m = _calc_resource(p, S)
            req[i] += np.array(res, np.int64)
            nz[i, 0] += c
            nz[i, 1] += m
            all_pods.append(p)
        pod_off.append(len(all_pods))
    pod_off = np.array(pod_off, np.int64)
    pods_b = build_podset(all_pods, S)
    sn = L.Snapshot()
    sn.n_nodes = N
    sn.n_scalars = S
    sn.node_valid, sn.alloc, sn.alloc_scalar_present = _p(valid, C.c_uint8), _p(alloc, C.c_int64), _p(present, C.c_uint32)
    sn.requested, sn.nonzero, sn.pod_off = _p(req, C.c_int64), _p(nz, C.c_int64), _p(pod_off, C.c_int64)
    sn.pods = pods_b.struct
    return Built(sn, [alloc, present, req, nz, valid, pod_off, pods_b])


def build_podgroups(groups, S=2):
    """groups: list of dict(exists, ns, min_member, min_resources (dict or None), denied, permitted)"""
    G = len(groups)
    K = 4 + S
    ex = np.array([1 if g.get("exists", True) else 0 for g in groups] or [0], np.uint8)
    ns = np.array([g.get("ns", 0) for g in groups] or [0], np.int32)
    mm = np.array([g["min_member"] for g in groups] or [0], np.int32)
    has = np.array([1 if g.get("min_resources") is not None else 0 for g in groups] or [0], np.uint8)
    den = np.array([1 if g.get("denied") else 0 for g in groups] or [0], np.uint8)
    per = np.array([1 if g.get("permitted") else 0 for g in groups] or [0], np.uint8)
    mr = np.zeros((max(G, 1), K), np.int64)
    mp = np.zeros(max(G, 1), np.uint32)
    for i, g in enumerate(groups):
        for k, v in (g.get("min_resources") or {}).items():
            col = 3 if k == "pods" else (4 + k[1] if isinstance(k, tuple) else _slot(k))
            mr[i, col] = v
            mp[i] |= 1 << col
    pg = L.Podgroups()
    pg.n_groups = G
    pg.n_scalars = S
    pg.exists, pg.ns, pg.min_member, pg.has_min_resources = _p(ex, C.c_uint8), _p(ns, C.c_int32), _p(mm, C.c_int32), _p(has, C.c_uint8)
    pg.min_res, pg.min_res_present, pg.denied, pg.permitted = _p(mr, C.c_int64), _p(mp, C.c_uint32), _p(den, C.c_uint8), _p(per, C.c_uint8)
    return Built(pg, [ex, ns, mm, has, den, per, mr, mp])


def build_quotas(quotas, n_ns, S=2):
    """quotas: {ns: dict(min={...}, max={...}, used={...})} with keys "cpu", "memory", ("scalar", s)"""
    W = 2 + S
    has = np.zeros(max(n_ns, 1), np.uint8)
    mn = np.zeros((max(n_ns, 1), W), np.int64)
    mx = np.zeros((max(n_ns, 1), W), np.int64)
    us = np.zeros((max(n_ns, 1), W), np.int64)
    up = np.zeros(max(n_ns, 1), np.uint32)

    def col(k):
        return 0 if k == "cpu" else 1 if k == "memory" else 2 + k[1]

    for k, q in quotas.items():
        has[k] = 1
        for r, v in q.get("min", {}).items():
            mn[k, col(r)] = v
        for r, v in q.get("max", {}).items():
            mx[k, col(r)] = v
        for r, v in q.get("used", {}).items():
            us[k, col(r)] = v
            if isinstance(r, tuple):
                up[k] |= 1 << r[1]
    qs = L.Quotas()
    qs.n_ns = n_ns
    qs.n_scalars = S
    qs.has_eq, qs.min, qs.max, qs.used, qs.used_scalar_present = _p(has, C.c_uint8), _p(mn, C.c_int64), _p(mx, C.c_int64), _p(us, C.c_int64), _p(up, C.c_uint32)
    return Built(qs, [has, mn, mx, us, up])

"""ctypes driver of the CPU oracle (oracle/libfgs_oracle.so). TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
The struct layouts are shared with the product's ctypes declarations because both sides read the
same flat snapshot format (include/fgs_types.h); no product code imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

import flex_gpu_scheduler_b200._lib as L

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libfgs_oracle.so")


class OracleOut(C.Structure):
    _fields_ = [
        ("status", L.p_u8), ("raw_flex", L.p_i64), ("raw_fit", L.p_i64), ("raw_alloc", L.p_i64),
        ("norm_flex", L.p_i64), ("norm_fit", L.p_i64), ("norm_alloc", L.p_i64), ("total", L.p_i64),
        ("feasible_count", L.p_i32), ("pod_err", L.p_u32), ("max_total", L.p_i64), ("tie_count", L.p_i32), ("first_max", L.p_i32),
    ]


_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        build()
    lib = C.CDLL(LIB_PATH)
    lib.fgs_oracle_evaluate.restype = C.c_int
    lib.fgs_oracle_evaluate.argtypes = [C.POINTER(L.Snapshot), C.POINTER(L.Podset), C.POINTER(L.PluginCfg), C.c_int64, C.c_int64,
                                        C.c_int, C.POINTER(OracleOut)]
    lib.fgs_oracle_reserve_index.restype = C.c_int
    lib.fgs_oracle_reserve_index.argtypes = [C.POINTER(L.Snapshot), C.POINTER(L.Podset), C.c_int64, C.c_int64]
    lib.fgs_oracle_gang_prefilter.restype = C.c_int
    lib.fgs_oracle_gang_prefilter.argtypes = [C.POINTER(L.Snapshot), C.POINTER(L.Podset), C.POINTER(L.Podgroups), C.c_int64, C.c_int64,
                                              L.p_u8]
    lib.fgs_oracle_gang_prefilter_listed.restype = C.c_int
    lib.fgs_oracle_gang_prefilter_listed.argtypes = [C.POINTER(L.Snapshot), C.POINTER(L.Podset), C.POINTER(L.Podgroups), C.c_int64,
                                                     C.c_int64, L.p_u8, L.p_u8]
    lib.fgs_oracle_check_cluster_resource_raw.restype = C.c_int
    lib.fgs_oracle_check_cluster_resource_raw.argtypes = [C.POINTER(L.Snapshot), L.p_i64, C.c_uint32, C.c_int32]
    lib.fgs_oracle_check_cluster_resource.restype = C.c_int
    lib.fgs_oracle_check_cluster_resource.argtypes = [C.POINTER(L.Snapshot), C.POINTER(L.Podgroups), C.c_int64]
    lib.fgs_oracle_gang_permit.restype = C.c_int
    lib.fgs_oracle_gang_permit.argtypes = [C.POINTER(L.Snapshot), C.POINTER(L.Podset), C.POINTER(L.Podgroups), C.c_int64, C.c_int64,
                                           L.p_u8, L.p_i32]
    lib.fgs_oracle_quota_prefilter.restype = C.c_int
    lib.fgs_oracle_quota_prefilter.argtypes = [C.POINTER(L.Podset), C.POINTER(L.Quotas), C.POINTER(L.Podset), L.p_i64, L.p_i64, C.c_int64,
                                               C.c_int64, L.p_u8]
    lib.fgs_oracle_num_feasible_nodes_to_find.restype = C.c_int32
    lib.fgs_oracle_num_feasible_nodes_to_find.argtypes = [C.c_int32, C.c_int32]
    lib.fgs_oracle_evaluate_sampled.restype = C.c_int
    lib.fgs_oracle_evaluate_sampled.argtypes = [C.POINTER(L.Snapshot), C.POINTER(L.Podset), C.POINTER(L.PluginCfg), C.c_int64, C.c_int64,
                                                C.c_int, C.POINTER(C.c_int64), L.p_i32, C.POINTER(OracleOut)]
    lib.fgs_oracle_evaluate_nominated.restype = C.c_int
    lib.fgs_oracle_evaluate_nominated.argtypes = [C.POINTER(L.Snapshot), C.POINTER(L.Podset), C.POINTER(L.PluginCfg), C.c_int64, C.c_int64,
                                                  C.c_int, C.POINTER(L.Podset), L.p_i32, L.p_i64, L.p_i64, C.POINTER(OracleOut)]
    lib.fgs_oracle_evaluate_masked.restype = C.c_int
    lib.fgs_oracle_evaluate_masked.argtypes = [C.POINTER(L.Snapshot), C.POINTER(L.Podset), C.POINTER(L.PluginCfg), C.c_int64, C.c_int64,
                                               C.c_int, L.p_u8, C.POINTER(OracleOut)]
    lib.fgs_oracle_pick_preemption_node.restype = C.c_int64
    lib.fgs_oracle_pick_preemption_node.argtypes = [C.POINTER(L.Snapshot), L.p_u8, L.p_u8, L.p_u8]
    lib.fgs_oracle_select_victims.restype = C.c_int
    lib.fgs_oracle_select_victims.argtypes = [C.POINTER(L.Snapshot), C.POINTER(L.Podset), C.c_int64, C.POINTER(L.PluginCfg),
                                              C.POINTER(L.Quotas), C.c_int, L.p_u8, L.p_u8, L.p_i32]
    lib.fgs_oracle_filter_reasons.restype = C.c_int
    lib.fgs_oracle_filter_reasons.argtypes = [C.POINTER(L.Snapshot), C.POINTER(L.Podset), C.POINTER(L.PluginCfg), C.c_int64,
                                              C.POINTER(C.c_uint16), C.POINTER(C.c_uint16)]
    lib.fgs_oracle_quota_reserve.restype = C.c_int
    lib.fgs_oracle_quota_reserve.argtypes = [C.POINTER(L.Podset), C.c_int64, C.c_int, L.p_i64, L.p_u32]
    _lib = lib
    return lib


class EvalResult:
    pass


def num_feasible_nodes_to_find(n_nodes, percentage=0):
    return int(load().fgs_oracle_num_feasible_nodes_to_find(n_nodes, percentage))


def evaluate(snapshot, pending, cfg, p0=0, p1=None, workers=1, planes=("status", "total"), per_plugin=False, sample=None,
             nominated=None, ext_mask=None):
    """Run the oracle on pods [p0,p1). Returns numpy arrays; planes are [p1-p0][N].
    sample=(percentage, start_index): the reference's node sampling at one worker; adds r.processed [P] and
    r.next_start (status 0xEE = node not visited).
    ext_mask: uint8 [P][N], 0 = a filter plugin outside this path rejected the node (status 0xE2): scoring, NormalizeScore
    and selectHost run over the remaining list."""
    lib = load()
    N = int(snapshot.contents.n_nodes)
    if p1 is None:
        p1 = int(pending.contents.n_pods)
    P = p1 - p0
    r = EvalResult()
    out = OracleOut()
    names = list(planes)
    if per_plugin:
        names += ["raw_flex", "raw_fit", "raw_alloc", "norm_flex", "norm_fit", "norm_alloc"]
    for nm in names:
        dt = np.uint8 if nm == "status" else np.int64
        a = np.zeros((P, N), dt)
        setattr(r, nm, a)
        setattr(out, nm, a.ctypes.data_as(L.p_u8 if nm == "status" else L.p_i64))
    r.feasible_count = np.zeros(P, np.int32)
    r.pod_err = np.zeros(P, np.uint32)
    r.max_total = np.zeros(P, np.int64)
    r.tie_count = np.zeros(P, np.int32)
    r.first_max = np.zeros(P, np.int32)
    out.feasible_count = r.feasible_count.ctypes.data_as(L.p_i32)
    out.pod_err = r.pod_err.ctypes.data_as(L.p_u32)
    out.max_total = r.max_total.ctypes.data_as(L.p_i64)
    out.tie_count = r.tie_count.ctypes.data_as(L.p_i32)
    out.first_max = r.first_max.ctypes.data_as(L.p_i32)
    if ext_mask is not None:
        em = np.ascontiguousarray(ext_mask, np.uint8)
        assert em.shape == (P, N)
        rc = lib.fgs_oracle_evaluate_masked(snapshot, pending, C.byref(cfg), p0, p1, workers, em.ctypes.data_as(L.p_u8), C.byref(out))
    elif nominated is not None:  # (podset pointer, node index array, uid array or None, pending uid array or None)
        nps, nnode, nuid, puid = nominated
        nnode = np.ascontiguousarray(nnode, np.int32)
        nu = np.ascontiguousarray(nuid, np.int64).ctypes.data_as(L.p_i64) if nuid is not None else None
        pu = np.ascontiguousarray(puid, np.int64).ctypes.data_as(L.p_i64) if puid is not None else None
        rc = lib.fgs_oracle_evaluate_nominated(snapshot, pending, C.byref(cfg), p0, p1, workers, nps, nnode.ctypes.data_as(L.p_i32), nu, pu,
                                               C.byref(out))
    elif sample is not None:
        start = C.c_int64(int(sample[1]))
        r.processed = np.zeros(max(P, 1), np.int32)
        rc = lib.fgs_oracle_evaluate_sampled(snapshot, pending, C.byref(cfg), p0, p1, int(sample[0]), C.byref(start),
                                             r.processed.ctypes.data_as(L.p_i32), C.byref(out))
        r.processed = r.processed[:P]
        r.next_start = int(start.value)
    else:
        rc = lib.fgs_oracle_evaluate(snapshot, pending, C.byref(cfg), p0, p1, workers, C.byref(out))
    if rc != 0:
        raise RuntimeError(f"fgs_oracle_evaluate rc={rc}")
    return r


def reserve_index(snapshot, pending, p, n):
    return load().fgs_oracle_reserve_index(snapshot, pending, p, n)


def gang_prefilter(snapshot, pending, podgroups, unlisted=None):
    P = int(pending.contents.n_pods)
    code = np.zeros(max(P, 1), np.uint8)
    ul = np.ascontiguousarray(unlisted, np.uint8).ctypes.data_as(L.p_u8) if unlisted is not None else None
    load().fgs_oracle_gang_prefilter_listed(snapshot, pending, podgroups, 0, P, code.ctypes.data_as(L.p_u8), ul)
    return code[:P]


def check_cluster_resource_raw(snapshot, request, keys, desired_pg):
    req = np.ascontiguousarray(request, np.int64)
    return load().fgs_oracle_check_cluster_resource_raw(snapshot, req.ctypes.data_as(L.p_i64), keys, desired_pg)


def check_cluster_resource(snapshot, podgroups, g):
    return load().fgs_oracle_check_cluster_resource(snapshot, podgroups, g)


def gang_permit(snapshot, pending, podgroups):
    P = int(pending.contents.n_pods)
    G = int(podgroups.contents.n_groups)
    st = np.zeros(max(P, 1), np.uint8)
    assigned = np.zeros(max(G, 1), np.int32)
    load().fgs_oracle_gang_permit(snapshot, pending, podgroups, 0, P, st.ctypes.data_as(L.p_u8), assigned.ctypes.data_as(L.p_i32))
    return st[:P], assigned[:G]


def quota_prefilter(pending, quotas, nominated=None, pending_uid=None, nominated_uid=None):
    P = int(pending.contents.n_pods)
    code = np.zeros(max(P, 1), np.uint8)
    pu = np.ascontiguousarray(pending_uid, np.int64).ctypes.data_as(L.p_i64) if pending_uid is not None else None
    nu = np.ascontiguousarray(nominated_uid, np.int64).ctypes.data_as(L.p_i64) if nominated_uid is not None else None
    load().fgs_oracle_quota_prefilter(pending, quotas, nominated, pu, nu, 0, P, code.ctypes.data_as(L.p_u8))
    return code[:P]


def quota_reserve(pods, p, used_row, used_present, op=1):
    """reserveResource (op=+1) / unreserveResource (op=-1) of pod p on one namespace's Used row (numpy int64 [2+S],
    modified in place); returns the new scalar-key mask."""
    assert used_row.dtype == np.int64 and used_row.flags["C_CONTIGUOUS"]
    up = np.array([used_present], np.uint32)
    load().fgs_oracle_quota_reserve(pods, p, op, used_row.ctypes.data_as(L.p_i64), up.ctypes.data_as(L.p_u32))
    return int(up[0])


def filter_reasons(snapshot, pending, cfg, p):
    """(flex reason ids [N], fit reason bit masks [N]) of pod p"""
    N = int(snapshot.contents.n_nodes)
    flex, fit = np.zeros(max(N, 1), np.uint16), np.zeros(max(N, 1), np.uint16)
    load().fgs_oracle_filter_reasons(snapshot, pending, C.byref(cfg), p, flex.ctypes.data_as(C.POINTER(C.c_uint16)),
                                     fit.ctypes.data_as(C.POINTER(C.c_uint16)))
    return flex[:N], fit[:N]


def select_victims(snapshot, pending, p, cfg, quotas=None, tracks_pods=True):
    """SelectVictimsOnNode for pod p on every node: (status [N], victim flags [pods in the snapshot], victims per node [N])"""
    N = int(snapshot.contents.n_nodes)
    T = int(snapshot.contents.pods.n_pods)
    st, vic, cnt = np.zeros(max(N, 1), np.uint8), np.zeros(max(T, 1), np.uint8), np.zeros(max(N, 1), np.int32)
    load().fgs_oracle_select_victims(snapshot, pending, p, C.byref(cfg), quotas, 1 if tracks_pods else 0, st.ctypes.data_as(L.p_u8),
                                     vic.ctypes.data_as(L.p_u8), cnt.ctypes.data_as(L.p_i32))
    return st[:N], vic[:T], cnt[:N]


def pick_preemption_node(snapshot, node_status, victim):
    """pickOneNodeForPreemption over the candidates of a dry run: the mask [N] of the nodes the reference may return
    (its last tie-break is Go map order)"""
    N = int(snapshot.contents.n_nodes)
    st = np.ascontiguousarray(node_status, np.uint8)
    vic = np.ascontiguousarray(victim if len(victim) else np.zeros(1, np.uint8), np.uint8)
    ties = np.zeros(max(N, 1), np.uint8)
    load().fgs_oracle_pick_preemption_node(snapshot, st.ctypes.data_as(L.p_u8), vic.ctypes.data_as(L.p_u8), ties.ctypes.data_as(L.p_u8))
    return ties[:N].astype(bool)

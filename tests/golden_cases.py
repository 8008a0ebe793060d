"""Turns the JSON golden vectors (tests/golden/*.json, transcribed from the reference's unit tests by
tests/golden/make_golden.py) into flat object snapshots. Shared by the oracle tests (CPU) and the
engine parity tests (GPU)."""
import json
import os

import builders as B

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    with open(os.path.join(GOLDEN, name + ".json")) as f:
        return json.load(f)


def _res(d):
    out = {}
    for k, v in d.items():
        if k.startswith("scalar:"):
            out[("scalar", int(k.split(":")[1]))] = v
        else:
            out[k] = v
    return out


# ---- allocatable ---------------------------------------------------------------------------------

def allocatable_case(case):
    """→ (snapshot Built, pending Built, resources list for plugin_cfg, mode)"""
    nodes = [B.node(alloc={"cpu": c, "memory": m, "pods": 110}) for c, m in case["nodes"]]
    snap = B.build_snapshot(nodes, S=2)
    # cpuAndMemory pod of the test (allocatable_test.go:80-83); its requests do not enter the score
    pend = B.build_podset([B.pod(containers=[B.ctr({"cpu": 1000, "memory": 1 << 30})])], S=2)
    return snap, pend, [tuple(r) for r in case["resources"]], case["mode"]


# ---- coscheduling ----------------------------------------------------------------------------------

def gang_prefilter_case(g, case):
    """→ (snapshot, pending, podgroups, unlisted mask, index of the evaluated pod)"""
    sn = g["snapshot"]
    n_nodes = sn["n_nodes"]
    per_node = [[] for _ in range(n_nodes)]
    for i in range(sn["n_existing_pods"]):
        per_node[i % n_nodes].append(B.pod(ns=0, pg=-1, node_name=True))
    snap = B.build_snapshot([B.node(alloc=_res(sn["node_alloc"]), pods=per_node[i]) for i in range(n_nodes)], S=2)
    groups = []
    for gi, gr in enumerate(g["groups"]):
        groups.append(dict(exists=gr["exists"], ns=gr["ns"], min_member=gr["min_member"],
                           min_resources=_res(gr["min_resources"]) if gr["min_resources"] else None,
                           denied=gi in case["denied"]))
    pgs = B.build_podgroups(groups, S=2)
    listed = [B.pod(ns=1, pg=pg) for pg in case["listed"]]
    ctrs = [B.ctr({"cpu": case["pod_cpu"]})] if case.get("pod_cpu") else []
    me = B.pod(ns=1, pg=case["pod_pg"], containers=ctrs)
    pend = B.build_podset(listed + [me], S=2)
    unlisted = [0] * len(listed) + [1]
    return snap, pend, pgs, unlisted, len(listed)


def gang_permit_case(g, case):
    groups = [dict(exists=gr["exists"], ns=gr["ns"], min_member=gr["min_member"]) for gr in g["groups"]]
    pgs = B.build_podgroups(groups, S=2)
    if case["with_snapshot"]:
        nodes = [B.node(alloc={"cpu": 1000, "pods": 20}, pods=[B.pod(ns=1, pg=1, node_name=True)])]
    else:
        nodes = []
    snap = B.build_snapshot(nodes, S=2)
    pend = B.build_podset([B.pod(ns=1, pg=case["pod_pg"])], S=2)
    return snap, pend, pgs


def check_cluster_resource_case(c, case):
    """→ (snapshot, request row [4+S], key mask, desired group id)"""
    nodes = [B.node(alloc=_res(c["node_alloc"]))]
    if case["with_assumed_pod"]:
        # the node-less NodeInfo of the fake lister, holding the pod; group id 1 = "/pg1-1", desired id 0 = "pg1-1"
        nodes.insert(0, B.node(alloc={}, valid=False, pods=[B.pod(ns=0, pg=1, containers=[B.ctr({"memory": 100})])]))
    snap = B.build_snapshot(nodes, S=2)
    req = [0] * 6
    keys = 0
    for k, v in case["request"].items():
        col = {"cpu": 0, "memory": 1, "ephemeral-storage": 2, "pods": 3}[k]
        req[col] = v
        keys |= 1 << col
    return snap, req, keys, 0


# ---- capacity scheduling -------------------------------------------------------------------------------

def capacity_case(case):
    """→ (pending, quotas) with S = 3: scalar slot 2 = nvidia.com/gpu"""
    S = 3
    n_ns = 3
    quotas = {int(k): {f: _res(v) for f, v in q.items()} for k, q in case["quotas"].items()}
    qs = B.build_quotas(quotas, n_ns, S=S)
    pods = [B.pod(ns=p["ns"], containers=[B.ctr({"memory": p["mem"], "cpu": 0, ("scalar", 2): 0})]) for p in case["pods"]]
    pend = B.build_podset(pods, S=S)
    return pend, qs

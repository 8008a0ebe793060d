"""CPU tests: the oracle (oracle/ref_walk.cpp) against the golden vectors transcribed from the
reference's own unit tests (tests/golden/make_golden.py cites every case)."""
import ctypes as C

import numpy as np
import pytest

import golden_cases as G
import flex_gpu_scheduler_b200 as fgs
from flex_gpu_scheduler_b200 import _lib as L

FW_CODE = {0: 0, 1: 2, 2: 2}  # oracle quota code → framework.Code (Success / Unschedulable)


@pytest.mark.parametrize("case", [c for c in G.load("allocatable") if "want_err" not in c], ids=lambda c: c["name"][:60])
def test_allocatable_scores(oracle, case):
    # pkg/noderesources/allocatable_test.go:112-219
    snap, pend, resources, mode = G.allocatable_case(case)
    cfg = fgs.plugin_cfg(filter_plugins=(), score_plugins=(("NodeResourcesAllocatable", 1),), alloc_mode=mode, alloc_resources=resources)
    r = oracle.evaluate(snap.ptr, pend.ptr, cfg, per_plugin=True)
    assert r.norm_alloc[0].tolist() == case["expected"]
    assert r.total[0].tolist() == case["expected"]
    assert int(r.pod_err[0]) == 0


def test_gang_prefilter(oracle):
    # pkg/coscheduling/core/core_test.go:42-176
    g = G.load("gang_prefilter")
    for case in g["cases"]:
        snap, pend, pgs, unlisted, me = G.gang_prefilter_case(g, case)
        code = oracle.gang_prefilter(snap.ptr, pend.ptr, pgs.ptr, unlisted=unlisted)
        assert (code[me] == 0) == case["expected"], case["name"]


def test_gang_permit(oracle):
    # pkg/coscheduling/core/core_test.go:178-234
    g = G.load("gang_permit")
    for case in g["cases"]:
        snap, pend, pgs = G.gang_permit_case(g, case)
        st, assigned = oracle.gang_permit(snap.ptr, pend.ptr, pgs.ptr)
        assert int(st[0]) == case["expected"], case["name"]


def test_check_cluster_resource(oracle):
    # pkg/coscheduling/core/core_test.go:303-355
    c = G.load("check_cluster_resource")
    for case in c["cases"]:
        snap, req, keys, desired = G.check_cluster_resource_case(c, case)
        ok = oracle.check_cluster_resource_raw(snap.ptr, req, keys, desired)
        assert bool(ok) == case["expected"], case["name"]


def test_capacity_prefilter(oracle):
    # pkg/capacityscheduling/capacity_scheduling_test.go:52-164
    for case in G.load("capacity_prefilter"):
        pend, qs = G.capacity_case(case)
        code = oracle.quota_prefilter(pend.ptr, qs.ptr)
        assert [FW_CODE[int(c)] for c in code] == case["expected"], case["name"]


def test_capacity_reject_reasons(oracle):
    # the two rejections of TestPreFilter come from different checks (capacity_scheduling.go:266 vs :270)
    cases = G.load("capacity_prefilter")
    pend, qs = G.capacity_case(cases[0])
    assert oracle.quota_prefilter(pend.ptr, qs.ptr).tolist() == [0, 1]
    pend, qs = G.capacity_case(cases[1])
    assert oracle.quota_prefilter(pend.ptr, qs.ptr).tolist() == [2]


def test_quota_reserve_golden(oracle):
    """TestReserveResource / TestUnReserveResource (pkg/capacityscheduling/elasticquota_test.go:27-129): Used ± the
    pods' requests for memory, cpu and an extended (scalar) resource."""
    S = 3
    import builders as B
    for case in G.load("quota_reserve"):
        pods = [B.pod([B.ctr({"cpu": pd["cpu"], "memory": pd["mem"], ("scalar", 2): pd["gpu"]})]) for pd in case["pods"]]
        pend = B.build_podset(pods, S)
        used = np.zeros(2 + S, np.int64)
        used[0], used[1], used[2 + 2] = case["before"]["cpu"], case["before"]["memory"], case["before"]["scalar:2"]
        present = 1 << 2
        for p in range(len(pods)):
            present = oracle.quota_reserve(pend.ptr, p, used, present, op=1 if case["op"] == "reserve" else -1)
        got = {"cpu": int(used[0]), "memory": int(used[1]), "scalar:2": int(used[2 + 2])}
        assert got == case["expected"], case["name"]
        assert present == 1 << 2


def test_num_feasible_nodes_to_find(oracle):
    """numFeasibleNodesToFind (vendor/k8s.io/kubernetes/pkg/scheduler/generic_scheduler.go:171-191). The vendored
    tree carries no _test.go files; these are the cases of upstream's TestNumFeasibleNodesToFind (kubernetes
    v1.22 pkg/scheduler/generic_scheduler_test.go) plus the K of the five BASELINE configs."""
    cases = [(10, 0, 10), (10, 40, 10), (1000, 0, 420), (1000, 40, 400), (6000, 0, 300), (6000, 40, 2400),
             (1000, 0, 420), (5000, 0, 500), (10000, 0, 500), (20000, 0, 1000), (50000, 0, 2500), (99, 10, 99), (5000, 100, 5000),
             (150, 10, 100)]
    for n, pct, want in cases:
        assert oracle.num_feasible_nodes_to_find(n, pct) == want, (n, pct)


def test_dry_run_preemption_golden(oracle):
    """TestDryRunPreemption (pkg/capacityscheduling/capacity_scheduling_test.go:166-282): the victims SelectVictimsOnNode
    picks for the in-namespace and the cross-namespace case."""
    import builders as B
    S = 3
    for case in G.load("dry_run_preemption"):
        def mk(d, **kw):
            return B.pod([B.ctr({"memory": d["mem"], "cpu": 0, ("scalar", 2): 0})], ns=d["ns"], priority=d["priority"], **kw)
        nodes = [B.node({"memory": case["node"]["memory"], "pods": case["node"]["pods"]}, [mk(d, node_name=True) for d in case["assigned"]])]
        snap = B.build_snapshot(nodes, S)
        pend = B.build_podset([mk(case["preemptor"])], S)
        n_ns = 1 + max(int(k) for k in case["quotas"])
        qs = B.build_quotas({int(k): v for k, v in case["quotas"].items()}, n_ns, S)
        cfg = fgs.plugin_cfg(filter_plugins=("NodeResourcesFit",), score_plugins=(("NodeResourcesFit", 1),))
        st, vic, cnt = oracle.select_victims(snap.ptr, pend.ptr, 0, cfg, qs.ptr, tracks_pods=False)
        assert int(st[0]) == L.FGS_PREEMPT_CANDIDATE, case["name"]
        got = [d["name"] for d, v in zip(case["assigned"], vic) if v]
        assert got == case["want_victims"] and int(cnt[0]) == len(got), (case["name"], got)


def test_pick_one_node_for_preemption(oracle):
    """pickOneNodeForPreemption (preemption.go:386-505) criterion by criterion; the vendored package carries no tests, so
    the expectations are the function's documented order: highest victim priority, sum of priorities (+2^31 each),
    number of victims, latest earliest-start-time of the highest-priority victims, then any of the rest"""
    import numpy as np
    import builders as B
    from flex_gpu_scheduler_b200 import _lib as L

    def run(per_node, status=None):
        """per_node: for every node a list of (priority, start, is_victim)"""
        nodes = [B.node({"cpu": 1000, "pods": 110}, [B.pod([B.ctr({"cpu": 1})], priority=pr, start=st, node_name=True) for pr, st, _ in pods])
                 for pods in per_node]
        snap = B.build_snapshot(nodes)
        vic = np.array([int(v) for pods in per_node for _, _, v in pods] or [0], np.uint8)
        st = np.array(status if status is not None else
                      [L.FGS_PREEMPT_CANDIDATE if any(v for _, _, v in pods) else L.FGS_PREEMPT_NO_VICTIMS for pods in per_node], np.uint8)
        return np.nonzero(oracle.pick_preemption_node(snap.ptr, st, vic))[0].tolist()

    V, K = True, False
    assert run([[(5, 0, V)]]) == [0]
    assert run([[(5, 0, K)], [(9, 0, K)]]) == []
    assert run([[(50, 0, V)], [(40, 0, V), (40, 0, V), (40, 0, V)], [(60, 0, V)]]) == [1]           # criterion 2 before 3 and 4
    assert run([[(50, 0, V), (50, 0, V)], [(50, 0, V), (10, 0, V)], [(50, 0, V), (20, 0, V)]]) == [1]
    assert run([[(-1, 0, V), (-1, 0, V)], [(-1, 0, V)]]) == [1]
    assert run([[(7, 0, V), (-2147483648, 0, V)], [(7, 0, V)]]) == [1]                              # equal sums: fewer victims
    assert run([[(5, 10, V)], [(5, 30, V)], [(5, 20, V)]]) == [1]
    assert run([[(5, 9, V), (5, 30, V)], [(5, 10, V), (5, 11, V)]]) == [1]
    assert run([[(5, 9, V), (4, 1, V)], [(5, 8, V), (4, 50, V)]]) == [0]                            # only the top-priority victims' times count
    assert run([[(5, 7, V)], [(3, 0, V)], [(5, 7, V)], [(3, 0, V)]]) == [1, 3]
    assert run([[(5, 7, V), (1, 0, K)], [(5, 7, V)]]) == [0, 1]                                     # pods that stay do not count
    # a node that is not a candidate is never picked, whatever its victim flags say
    assert run([[(1, 0, V)], [(9, 0, V)]], status=[L.FGS_PREEMPT_STILL_UNFIT, L.FGS_PREEMPT_CANDIDATE]) == [1]

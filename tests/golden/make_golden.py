"""Writes the golden-vector fixtures under tests/golden/.

The reference is Go and cannot be run in this image (no Go toolchain), so these vectors are
TRANSCRIBED by hand from the table-driven unit tests of the reference; every case cites its
source lines (paths relative to /root/reference). Run ``python tests/golden/make_golden.py`` to
regenerate the JSON files. Quantities are already converted the way the Go fixtures convert
them (cpu "1" → 1000 milli, resource.MustParse("1Gi") → 1073741824, …).

Resource keys in the JSON: "cpu" (milli), "memory", "pods", "scalar:<slot>".
"""
import json
import os

HERE = os.path.dirname(os.path.abspath(__file__))
Gi = 1 << 30

# ------------------------------------------------------------------------------------------------
# pkg/noderesources/allocatable_test.go:40-236 (TestNodeResourcesAllocatable)
# makeNodeInfo(name, milliCPU, memory) :303-319. The plugin is exercised Score → NormalizeScore on
# all listed nodes (:281-295); the pod's own requests never enter the score (allocatable.go:61-72).
default_set = [["cpu", 1 << 20], ["memory", 1]]          # :88-91
cpu_set = [["cpu", 1 << 30], ["memory", 1]]              # :94-97
allocatable = [
    dict(name="nothing scheduled, nothing requested", src="allocatable_test.go:112-118",
         nodes=[[4000, 10000], [4000, 10000]], resources=default_set, mode="Least", expected=[0, 0]),
    dict(name="nothing scheduled, resources requested, differently sized machines, least mode", src=":119-125",
         nodes=[[4000, 10000], [6000, 10000]], resources=default_set, mode="Least", expected=[100, 0]),
    dict(name="nothing scheduled, resources requested, differently sized machines, most mode", src=":126-132",
         nodes=[[4000, 10000], [6000, 10000]], resources=default_set, mode="Most", expected=[0, 100]),
    dict(name="no resources requested, pods scheduled", src=":133-145",
         nodes=[[4000, 10000], [4000, 10000]], resources=default_set, mode="Least", expected=[0, 0]),
    dict(name="no resources requested, pods scheduled with resources", src=":146-155",
         nodes=[[10000, 20000], [10000, 20000]], resources=default_set, mode="Least", expected=[0, 0]),
    dict(name="resources requested, pods scheduled with resources", src=":156-165",
         nodes=[[10000, 20000], [10000, 20000]], resources=default_set, mode="Least", expected=[0, 0]),
    dict(name="resources requested with more than the node, differently sized machines, least mode", src=":166-172",
         nodes=[[4000, 1000], [5000, 1000]], resources=default_set, mode="Least", expected=[100, 0]),
    dict(name="resources requested with more than the node, differently sized machines, most mode", src=":173-179",
         nodes=[[4000, 1000], [5000, 1000]], resources=default_set, mode="Most", expected=[0, 100]),
    dict(name="nothing scheduled, resources requested, differently sized machines, cpu weighted, most mode", src=":180-186",
         nodes=[[1000, 2000], [1005, 1000]], resources=cpu_set, mode="Least", expected=[100, 0]),
    dict(name="nothing scheduled, resources requested, differently sized machines, cpu weighted, least mode", src=":187-193",
         nodes=[[1000, 2000], [1005, 1000]], resources=cpu_set, mode="Most", expected=[0, 100]),
    dict(name="nothing scheduled, resources requested, 3 differently sized machines, least mode", src=":194-206",
         nodes=[[1000, 1000 << 20], [2000, 2000 << 20], [3000, 3000 << 20]], resources=default_set, mode="Least",
         expected=[100, 50, 0]),
    dict(name="nothing scheduled, resources requested, 3 differently sized machines, most mode", src=":207-219",
         nodes=[[1000, 1000 << 20], [2000, 2000 << 20], [3000, 3000 << 20]], resources=default_set, mode="Most",
         expected=[0, 50, 100]),
    dict(name="resource with negative weight", src=":220-227", nodes=[[4000, 10000]], resources=[["memory", -1], ["cpu", 1]],
         mode="Least", want_err="should be a positive value, got -1"),
    dict(name="resource with zero weight", src=":228-235", nodes=[[4000, 10000]], resources=[["memory", 1], ["cpu", 0]],
         mode="Least", want_err="should be a positive value, got 0"),
]

# ------------------------------------------------------------------------------------------------
# pkg/coscheduling/core/core_test.go:42-176 (TestPreFilter). Snapshot = MakeNodesAndPods({"test":"a"},
# 60 pods, 30 nodes) (test/util/utils.go:47-75): nodes cpu "1", pods "20"; pod i on node i%30, no
# containers, no pod-group label, default namespace. Groups (:51-54): pg, pg1 (min 2), pg2 (min 2,
# MinResources cpu 4), pg3 (min 2, MinResources cpu 40), all in ns1. `pods` are what the test puts
# into the pod informer; the evaluated pod itself is NOT listed (:159-164).
# Namespace ids: 0 = default, 1 = ns1. Group ids: 0 pg, 1 pg1, 2 pg2, 3 pg3, 4 pg-notexisting.
gang_groups = [
    dict(name="pg", ns=1, min_member=2, min_resources=None, exists=True),
    dict(name="pg1", ns=1, min_member=2, min_resources=None, exists=True),
    dict(name="pg2", ns=1, min_member=2, min_resources={"cpu": 4000}, exists=True),
    dict(name="pg3", ns=1, min_member=2, min_resources={"cpu": 40000}, exists=True),
    dict(name="pg-notexisting", ns=1, min_member=0, min_resources=None, exists=False),
]
gang_prefilter = dict(
    snapshot=dict(n_nodes=30, node_alloc={"cpu": 1000, "pods": 20}, n_existing_pods=60),
    groups=gang_groups,
    cases=[
        dict(name="pod does not belong to any pg", src="core_test.go:70-79", pod_pg=-1, listed=[1, 2], denied=[], expected=True),
        dict(name="pg was previously denied", src=":80-85", pod_pg=1, listed=[], denied=[1], expected=False),
        dict(name="pod belongs to a non-existing pg", src=":86-91", pod_pg=4, listed=[], denied=[], expected=True),
        dict(name="pod count less than minMember", src=":92-100", pod_pg=1, listed=[2], denied=[], expected=False),
        dict(name="pod count equal minMember", src=":101-110", pod_pg=1, listed=[1, 1], denied=[], expected=True),
        dict(name="pod count more minMember", src=":111-121", pod_pg=1, listed=[1, 1, 1], denied=[], expected=True),
        dict(name="cluster resource enough, min Resource", src=":122-132", pod_pg=2, pod_cpu=1000, listed=[2, 2], denied=[],
             expected=True),
        dict(name="cluster resource not enough, min Resource", src=":133-143", pod_pg=3, pod_cpu=20000, listed=[3, 3], denied=[],
             expected=False),
        dict(name="cluster resource enough not required", src=":144-153", pod_pg=1, listed=[1, 1], denied=[], expected=True),
    ],
)

# core_test.go:178-234 (TestPermit). Snapshot (:191-194): one node, one pod labelled pg1, namespace ns1,
# Spec.NodeName set. Status: 0 Success, 1 Wait, 2 PodGroupNotSpecified, 3 PodGroupNotFound.
gang_permit = dict(
    groups=[dict(name="pg", ns=1, min_member=2, exists=True), dict(name="pg1", ns=1, min_member=2, exists=True),
            dict(name="pg-noexist", ns=1, min_member=0, exists=False)],
    cases=[
        dict(name="pod does not belong to any pg, allow", src="core_test.go:203-207", pod_pg=-1, with_snapshot=False, expected=2),
        dict(name="pod belongs to a non-existing pg", src=":208-212", pod_pg=2, with_snapshot=False, expected=3),
        dict(name="pod belongs to a pg that doesn't have enough pods", src=":213-218", pod_pg=1, with_snapshot=False, expected=1),
        dict(name="pod belongs to a pg that has enough pods", src=":219-224", pod_pg=1, with_snapshot=True, expected=0),
    ],
)

# core_test.go:303-355 (TestCheckClusterResource). Node "fake-node": memory 300 (+ pods 32 from
# st.MakeNode().Capacity, vendor/.../scheduler/testing/wrappers.go:483-492). Case 3's pod (memory 100, label
# pg1-1, no node name) sits in a node-less NodeInfo that the walk skips (test/util/fake.go:60-68,
# core.go:324-326); its full group name "/pg1-1" differs from the desired "pg1-1".
check_cluster_resource = dict(
    node_alloc={"memory": 300, "pods": 32},
    cases=[
        dict(name="Cluster resource enough", src="core_test.go:321-328", request={"memory": 10}, with_assumed_pod=False, expected=True),
        dict(name="Cluster resource not enough", src=":329-336", request={"memory": 1000}, with_assumed_pod=False, expected=False),
        dict(name="Cluster resource enough, some resources of the pods belonging to the group have been included", src=":337-345",
             request={"memory": 250}, with_assumed_pod=True, expected=True),
    ],
)

# ------------------------------------------------------------------------------------------------
# pkg/capacityscheduling/capacity_scheduling_test.go:52-164 (TestPreFilter). makePod (:371-384) always
# puts memory, cpu (0) and nvidia.com/gpu (0) into the requests → scalar slot 2 carries a present key
# with value 0. Codes: 0 Success, 2 Unschedulable (framework.Code).
capacity_prefilter = [
    dict(name="pod subjects to ElasticQuota", src="capacity_scheduling_test.go:66-90",
         quotas={"1": dict(min={"memory": 1000}, max={"memory": 2000}, used={"memory": 300})},
         pods=[dict(ns=1, mem=500), dict(ns=1, mem=1800)], expected=[0, 2]),
    dict(name="the sum of used is bigger than the sum of min", src=":91-124",
         quotas={"1": dict(min={"memory": 1000}, max={"memory": 2000}, used={"memory": 1800}),
                 "2": dict(min={"memory": 1000}, max={"memory": 2000}, used={"memory": 200})},
         pods=[dict(ns=2, mem=500)], expected=[2]),
]

# pkg/capacityscheduling/elasticquota_test.go:27-129 (TestReserveResource / TestUnReserveResource):
# Used ± computePodResourceRequest(pod) for memory, cpu and scalars (elasticquota.go:74-88).
quota_reserve = [
    dict(name="ElasticQuotaInfo ReserveResource", src="elasticquota_test.go:34-62", op="reserve",
         before={"cpu": 1000, "memory": 200, "scalar:2": 2},
         pods=[dict(mem=50, cpu=1000, gpu=1), dict(mem=100, cpu=2000, gpu=0), dict(mem=0, cpu=0, gpu=2)],
         expected={"cpu": 4000, "memory": 350, "scalar:2": 5}),
    dict(name="ElasticQuotaInfo UnReserveResource", src=":84-112", op="unreserve",
         before={"cpu": 4000, "memory": 200, "scalar:2": 5},
         pods=[dict(mem=50, cpu=1000, gpu=1), dict(mem=100, cpu=2000, gpu=0), dict(mem=0, cpu=0, gpu=2)],
         expected={"cpu": 1000, "memory": 50, "scalar:2": 2}),
]


# pkg/capacityscheduling/capacity_scheduling_test.go:166-282 (TestDryRunPreemption): one node with memory 150 (and the
# wrapper's 32 pods, vendor/…/scheduler/testing/wrappers.go:483-492), three assigned pods of memory 50, a preemptor of
# memory 50 and highPriority (1000; midPriority = 100, :49). The fixtures' ElasticQuotaInfos carry no pod sets and the
# test framework registers no CapacityScheduling PreFilter extensions, so Used does not move while pods are removed.
dry_run_preemption = [
    dict(name="in-namespace preemption", src="capacity_scheduling_test.go:178-230",
         node=dict(memory=150, pods=32), preemptor=dict(ns=1, mem=50, priority=1000),
         assigned=[dict(name="t1-p1", ns=1, mem=50, priority=100), dict(name="t1-p2", ns=2, mem=50, priority=100),
                   dict(name="t1-p3", ns=2, mem=50, priority=100)],
         quotas={"1": dict(min={"memory": 50}, max={"memory": 200}, used={"memory": 50}),
                 "2": dict(min={"memory": 200}, max={"memory": 200}, used={"memory": 100})},
         want_victims=["t1-p1"]),
    dict(name="cross-namespace preemption", src=":231-282",
         node=dict(memory=150, pods=32), preemptor=dict(ns=1, mem=50, priority=1000),
         assigned=[dict(name="t1-p1", ns=1, mem=50, priority=100), dict(name="t1-p2", ns=2, mem=50, priority=1000),
                   dict(name="t1-p3", ns=2, mem=50, priority=100)],
         quotas={"1": dict(min={"memory": 150}, max={"memory": 200}, used={"memory": 50}),
                 "2": dict(min={"memory": 50}, max={"memory": 200}, used={"memory": 100})},
         want_victims=["t1-p3"]),
]


def main():
    out = dict(allocatable=allocatable, gang_prefilter=gang_prefilter, gang_permit=gang_permit,
               check_cluster_resource=check_cluster_resource, capacity_prefilter=capacity_prefilter, quota_reserve=quota_reserve,
               dry_run_preemption=dry_run_preemption)
    for name, obj in out.items():
        with open(os.path.join(HERE, name + ".json"), "w") as f:
            json.dump(obj, f, indent=1, sort_keys=True)
            f.write("\n")


if __name__ == "__main__":
    main()

"""CPU tests: hand-derived known-answer cases for the parts of the path the reference has NO tests for
(pkg/flexgpu has no _test.go; the vendored in-tree NodeResourcesFit / DefaultNormalizeScore / selectHost
ship without tests). Each expectation below is worked out by hand from the cited Go lines; they pin the
oracle where no reference vector exists ("parity unpinned" in DESIGN.md refers to exactly this gap)."""
import numpy as np

import builders as B
import flex_gpu_scheduler_b200 as fgs
from flex_gpu_scheduler_b200 import _lib as L

FLEX_ONLY = dict(filter_plugins=("FlexGPU",), score_plugins=(("FlexGPU", 1),))


def flex_node(gpus, mem_alloc, used=(), mono=(), extra_pods=(), cpu=64000, mem=1 << 40, has_gpu=True, has_mem=True):
    pods = [B.pod(containers=[B.ctr({("scalar", 1): u}, mem_limit=u)], node_name=True, index=i) for i, u in enumerate(used) if u]
    pods += [B.pod(containers=[B.ctr({("scalar", 0): 1}, gpu_limit=1)], node_name=True, index=i) for i in mono]
    pods += list(extra_pods)
    alloc = {"cpu": cpu, "memory": mem, "pods": 110}
    if has_gpu:
        alloc[("scalar", 0)] = gpus
    if has_mem:
        alloc[("scalar", 1)] = mem_alloc
    return B.node(alloc=alloc, pods=pods)


def mem_pod(limit, cpu=100):
    return B.pod(containers=[B.ctr({"cpu": cpu}, mem_limit=limit)])


def gpu_pod(limit=1, cpu=100):
    return B.pod(containers=[B.ctr({"cpu": cpu}, gpu_limit=limit)])


def test_filter_steps(oracle):
    # flex_gpu.go:41-119, one node per branch
    nodes = [
        flex_node(2, 2000),                                  # 0: empty → fits
        flex_node(2, 2000, has_mem=False),                   # 1: memory key missing → UnschedulableAndUnresolvable (:70-74)
        flex_node(2, 2000, has_gpu=False),                   # 2: gpu key missing → unresolvable (:64-68)
        flex_node(2, 2000, used=[900, 900]),                 # 3: sum 1800+300 > 2000 → Unschedulable (:95-98)
        flex_node(2, 2000, used=[800, 800]),                 # 4: sum 1900 ok, no device has 300 free (1000 < 1100) → Unschedulable (:109-116)
        flex_node(2, 2000, used=[800, 600]),                 # 5: device 1: 1000 >= 900 → fits
        flex_node(2, 2000, used=[0, 0], mono=[0, 1]),        # 6: both devices monopolised; gpu sum 2 <= 2 passes → no device fits
        flex_node(2, 2000, mono=[0, 1, 1]),                  # 7: gpu limit sum 3 > 2 → Unschedulable for ANY flex pod (:90-93)
    ]
    snap = B.build_snapshot(nodes)
    pend = B.build_podset([mem_pod(300), gpu_pod(1), B.pod(containers=[B.ctr({"cpu": 100})]),
                           B.pod(containers=[B.ctr({"cpu": 100}, gpu_limit=1, mem_limit=10)])])
    r = oracle.evaluate(snap.ptr, pend.ptr, fgs.plugin_cfg(**FLEX_ONLY), per_plugin=True)
    U, UU = 0x12, 0x13  # FlexGPU failed with Unschedulable / UnschedulableAndUnresolvable
    assert r.status[0].tolist() == [0, UU, UU, U, U, 0, U, U]
    # whole-GPU pod: only the two sums are checked (no per-device test): node 6 has sum 2+1 > 2
    assert r.status[1].tolist() == [0, UU, UU, 0, 0, 0, U, U]
    # no flex limit → Success everywhere (:54-56); Score then runs NewGPUNode on nodes 1,2?? node 1 HAS the gpu key
    # (count 2) so no panic there; node 2 has gpu count 0 → divide by zero (gpu_node.go:55)
    assert r.status[2].tolist() == [0] * 8
    assert int(r.pod_err[2]) == L.FGS_PODERR_REF_PANIC
    # both limits → unresolvable on every node (:58-61)
    assert r.status[3].tolist() == [UU] * 8
    assert r.feasible_count.tolist() == [2, 4, 8, 0]


def test_scores_and_reverse_normalisation(oracle):
    # GPUScore = idle, non-monopolised devices; MemScore = Σ(memEach − used) (gpu_node.go:179-199);
    # DefaultNormalizeScore(100, reverse) (normalize_score.go:26-54): 100 − 100·s/max, integer division
    nodes = [flex_node(4, 4000, used=[100, 0, 0, 0]), flex_node(4, 4000, used=[0, 0, 0, 0], mono=[3]), flex_node(4, 4000),
             flex_node(4, 4003, used=[333, 0, 0, 0])]
    snap = B.build_snapshot(nodes)
    pend = B.build_podset([mem_pod(10), gpu_pod(1)])
    r = oracle.evaluate(snap.ptr, pend.ptr, fgs.plugin_cfg(**FLEX_ONLY), per_plugin=True)
    assert r.raw_flex[0].tolist() == [3900, 4000, 4000, 4000 - 333]  # node 3: memEach = 4003/4 = 1000
    assert r.norm_flex[0].tolist() == [100 - 390000 // 4000, 0, 0, 100 - 366700 // 4000]  # [3, 0, 0, 9]
    assert r.raw_flex[1].tolist() == [3, 3, 4, 3]
    assert r.norm_flex[1].tolist() == [25, 25, 0, 25]
    # selectHost picks among the maximum: fractional pod → node 3 alone; whole-GPU pod → tie of nodes 0,1,3
    assert r.first_max.tolist() == [3, 0]
    assert r.tie_count.tolist() == [1, 3]


def test_all_zero_scores_become_100(oracle):
    # maxCount == 0 with reverse → every score = 100 (normalize_score.go:34-41): pods without flex limits
    snap = B.build_snapshot([flex_node(2, 2000), flex_node(2, 2000, used=[5, 0])])
    pend = B.build_podset([B.pod(containers=[B.ctr({"cpu": 100})])])
    r = oracle.evaluate(snap.ptr, pend.ptr, fgs.plugin_cfg(**FLEX_ONLY), per_plugin=True)
    assert r.raw_flex[0].tolist() == [0, 0]
    assert r.norm_flex[0].tolist() == [100, 100]


def test_annotation_quirks(oracle):
    # gpu_node.go:85-120: pods without limits are skipped; a missing / unparsable annotation skips the pod for the
    # device table but its limit still counts in nodeResourceLimitSum (flex_gpu.go:133-140)
    no_annot = B.pod(containers=[B.ctr({("scalar", 1): 900}, mem_limit=900)], node_name=True, index=None)
    nodes = [flex_node(2, 2000, extra_pods=[no_annot]),     # device table empty, sum 900
             flex_node(2, 2000, extra_pods=[no_annot, no_annot])]  # sum 1800
    snap = B.build_snapshot(nodes)
    pend = B.build_podset([mem_pod(300), mem_pod(1000)])
    r = oracle.evaluate(snap.ptr, pend.ptr, fgs.plugin_cfg(**FLEX_ONLY), per_plugin=True)
    assert r.status[0].tolist() == [0, 0x12]        # 1800 + 300 > 2000
    assert r.status[1].tolist() == [0, 0x12]        # 900 + 1000 <= 2000 and an (apparently) empty device of 1000 fits
    assert r.raw_flex[0, 0] == 2000                 # MemScore ignores the un-annotated pod


def test_reference_panics(oracle):
    # index annotation outside [0, gpuCount) → index out of range in constructGPUs (gpu_node.go:110,115);
    # the panic happens only when FlexGPU.Filter gets as far as NewGPUNode (flex_gpu.go:101)
    bad = B.pod(containers=[B.ctr({("scalar", 1): 100}, mem_limit=100)], node_name=True, index=5)
    nodes = [flex_node(2, 2000, extra_pods=[bad]), flex_node(2, 2000)]
    snap = B.build_snapshot(nodes)
    pend = B.build_podset([mem_pod(300), mem_pod(5000), B.pod(containers=[B.ctr({"cpu": 1})])])
    r = oracle.evaluate(snap.ptr, pend.ptr, fgs.plugin_cfg(**FLEX_ONLY))
    assert r.status[0].tolist() == [0xF1, 0]
    assert int(r.pod_err[0]) == L.FGS_PODERR_REF_PANIC
    assert r.status[1].tolist() == [0x12, 0x12]     # rejected at the sum check before NewGPUNode: no panic
    assert int(r.pod_err[1]) == 0
    assert r.status[2].tolist() == [0, 0] and int(r.pod_err[2]) == L.FGS_PODERR_REF_PANIC  # Score calls NewGPUNode (:148)
    # Fit first and failing → FlexGPU.Filter never runs on that node (runtime/framework.go:674-689)
    cfg = fgs.plugin_cfg(filter_plugins=("NodeResourcesFit", "FlexGPU"), score_plugins=(("FlexGPU", 1),))
    big = B.build_podset([B.pod(containers=[B.ctr({"cpu": 10 ** 9}, mem_limit=300)])])
    r = oracle.evaluate(snap.ptr, big.ptr, cfg)
    assert r.status[0].tolist() == [0x22, 0x22] and int(r.pod_err[0]) == 0


def test_reserve_first_fit_and_aliasing(oracle):
    # MemAssumeFitIndexes (gpu_node.go:122-161): `remain := u.memory; remain.Sub(*assumed)` shrinks the SHARED
    # per-device capacity after every fit, and the sort compares a cell with itself → order unchanged.
    # Devices of 1000, used [700, 0, 200], limit 100: device 0 fits (1000 >= 800) → shared capacity becomes 200;
    # device 1: 200 >= 100 fits → capacity 100; device 2: 100 >= 300 fails. indexes = [0, 1] → Reserve writes "0"
    # (first fit), although best-fit bin-packing was intended.
    snap = B.build_snapshot([flex_node(3, 3000, used=[700, 0, 200]), flex_node(3, 3000, used=[950, 0, 200]),
                             flex_node(2, 2000, mono=[0]), flex_node(2, 2000, used=[10, 0], mono=[1])])
    pend = B.build_podset([mem_pod(100), gpu_pod(1), B.pod(containers=[B.ctr({"cpu": 1})])])
    assert oracle.reserve_index(snap.ptr, pend.ptr, 0, 0) == 0
    assert oracle.reserve_index(snap.ptr, pend.ptr, 0, 1) == 1      # device 0: 1000 < 1050
    assert oracle.reserve_index(snap.ptr, pend.ptr, 1, 2) == 1      # whole GPU: first idle non-monopolised device
    assert oracle.reserve_index(snap.ptr, pend.ptr, 1, 3) == -1     # device 0 used, device 1 monopolised
    assert oracle.reserve_index(snap.ptr, pend.ptr, 2, 0) == -2     # nothing to reserve


def test_fit_filter_and_least_most(oracle):
    # fit.go:256-329 and least_allocated.go:46-57 / most_allocated.go:49-61 with non-zero defaults
    # (pod_resources.go:35-37): a container without cpu/memory requests counts 100m / 200Mi for SCORING only.
    Mi = 1 << 20
    existing = B.pod(containers=[B.ctr({"cpu": 1000, "memory": 1000 * Mi})], node_name=True)
    nodes = [B.node(alloc={"cpu": 4000, "memory": 4000 * Mi, "pods": 2}, pods=[existing]),
             B.node(alloc={"cpu": 4000, "memory": 4000 * Mi, "pods": 1}, pods=[existing]),     # too many pods
             B.node(alloc={"cpu": 1500, "memory": 4000 * Mi, "pods": 10}, pods=[existing]),    # 1000 > 1500-1000 cpu
             B.node(alloc={"cpu": 0, "memory": 4000 * Mi, "pods": 10})]                        # cpu capacity 0
    snap = B.build_snapshot(nodes)
    pend = B.build_podset([B.pod(containers=[B.ctr({"cpu": 1000, "memory": 500 * Mi})]),
                           B.pod(containers=[B.ctr({})])])  # requests nothing: only the pod count is checked (:270-275)
    cfg = fgs.plugin_cfg(filter_plugins=("NodeResourcesFit",), score_plugins=(("NodeResourcesFit", 1),))
    r = oracle.evaluate(snap.ptr, pend.ptr, cfg, per_plugin=True)
    assert r.status[0].tolist() == [0, 0x22, 0x22, 0x22]
    assert r.status[1].tolist() == [0, 0x22, 0, 0]
    # pod 0 on node 0: cpu (4000-2000)*100/4000 = 50, memory (4000-1500)*100/4000 = 62 → (50+62)/2 = 56
    assert r.raw_fit[0, 0] == 56
    # pod 1 (defaults 100m / 200Mi): node 0: cpu (4000-1100)*100/4000 = 72, mem (4000-1200)*100/4000 = 70 → 71
    #   node 2: cpu (1500-1100)*100/1500 = 26, mem 70 → 48;  node 3: cpu capacity 0 is left out of the average
    #   (resource_allocation.go:64-67): mem (4000-200)*100/4000 = 95 → 95
    assert r.raw_fit[1].tolist() == [71, 0, 48, 95]
    most = fgs.plugin_cfg(filter_plugins=("NodeResourcesFit",), score_plugins=(("NodeResourcesFit", 1),), fit_strategy="MostAllocated",
                          fit_resources=(("cpu", 3), ("memory", 1)))
    r = oracle.evaluate(snap.ptr, pend.ptr, most, per_plugin=True)
    # pod 1: node 0: cpu 1100*100/4000 = 27, mem 1200*100/4000 = 30 → (27*3+30)/4 = 27; node 2: cpu 73, mem 30 → 62; node 3: mem 5 → 5
    assert r.raw_fit[1].tolist() == [27, 0, 62, 5]


def test_synthetic_configs_exercise_every_branch(oracle):
    s = fgs.config_synth(2, n_nodes=400, n_pending=300, pm_quirk_nodes=150, pm_panic_nodes=60, pm_nongpu_nodes=80)
    r = oracle.evaluate(s.snapshot, s.pending, fgs.config_plugins(2), workers=2)
    codes = set(np.unique(r.status).tolist())
    assert {0x00, 0x12, 0x13, 0x22, 0xF1} <= codes
    assert (r.pod_err & L.FGS_PODERR_REF_PANIC).any()
    assert r.feasible_count.max() > 100 and r.feasible_count.min() == 0

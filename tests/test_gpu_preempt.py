"""Batched dry-run preemption (fgs_select_victims): preemptor.SelectVictimsOnNode
(pkg/capacityscheduling/capacity_scheduling.go:465-644) for one pod on every node, against the oracle's restatement
and the reference's TestDryRunPreemption cases."""
import ctypes as C

import numpy as np
import pytest

import builders as B
import golden_cases as G
import flex_gpu_scheduler_b200 as fgs
from flex_gpu_scheduler_b200 import _lib as L

pytestmark = pytest.mark.gpu
GiB = 1 << 30


def _engine(snapshot_ptr, cfg, S=2):
    eng = fgs.Engine(device=0, n_scalars=S)
    eng.set_plugins(cfg)
    eng.upload_nodes(fgs.Packer(S).nodes(snapshot_ptr))
    return eng


def test_reference_dry_run_preemption_cases(gpu_ok, oracle):
    S = 3
    for case in G.load("dry_run_preemption"):
        def mk(d, **kw):
            return B.pod([B.ctr({"memory": d["mem"], "cpu": 0, ("scalar", 2): 0})], ns=d["ns"], priority=d["priority"], **kw)
        nodes = [B.node({"memory": case["node"]["memory"], "pods": case["node"]["pods"]}, [mk(d, node_name=True) for d in case["assigned"]])]
        snap = B.build_snapshot(nodes, S)
        pend = B.build_podset([mk(case["preemptor"])], S)
        qs = B.build_quotas({int(k): v for k, v in case["quotas"].items()}, 1 + max(int(k) for k in case["quotas"]), S)
        cfg = fgs.plugin_cfg(filter_plugins=("NodeResourcesFit",), score_plugins=(("NodeResourcesFit", 1),))
        eng = _engine(snap.ptr, cfg, S)
        st, vic, cnt = eng.select_victims(snap.ptr, pend.ptr, 0, qs.ptr, tracks_pods=False)
        assert int(st[0]) == L.FGS_PREEMPT_CANDIDATE, case["name"]
        assert [d["name"] for d, v in zip(case["assigned"], vic) if v] == case["want_victims"], case["name"]
        ref = oracle.select_victims(snap.ptr, pend.ptr, 0, cfg, qs.ptr, tracks_pods=False)
        assert np.array_equal(st, ref[0]) and np.array_equal(vic, ref[1]) and np.array_equal(cnt, ref[2])
        eng.close()


@pytest.mark.parametrize("tracks", [True, False])
@pytest.mark.parametrize("with_quotas", [True, False])
def test_synthetic_cluster_matches_the_oracle(gpu_ok, oracle, tracks, with_quotas):
    s = fgs.config_synth(4, n_nodes=500, n_pending=60, n_namespaces=9, max_existing_pods=20, pm_pods_without_eq=200, pm_invalid_nodes=10,
                         pm_panic_nodes=10)
    cfg = fgs.config_plugins(4)
    eng = _engine(s.snapshot, cfg)
    q = s.quotas if with_quotas else None
    seen, picks = set(), []
    for p in range(0, s.n_pending, 2):
        st, vic, cnt, picked = eng.select_victims(s.snapshot, s.pending, p, q, tracks_pods=tracks, pick=True)
        rst, rvic, rcnt = oracle.select_victims(s.snapshot, s.pending, p, cfg, q, tracks_pods=tracks)
        assert np.array_equal(st, rst), (p, np.nonzero(st != rst)[0][:5], st[st != rst][:5], rst[st != rst][:5])
        assert np.array_equal(vic, rvic), p
        assert np.array_equal(cnt, rcnt), p
        # pickOneNodeForPreemption: the reference may return any node of the oracle's tie set; the engine the lowest
        ties = np.nonzero(oracle.pick_preemption_node(s.snapshot, rst, rvic))[0]
        assert picked == (int(ties[0]) if len(ties) else -1), (p, picked, ties[:5])
        picks.append(picked)
        seen |= set(int(x) for x in st)
    assert {L.FGS_PREEMPT_CANDIDATE, L.FGS_PREEMPT_NO_VICTIMS, L.FGS_PREEMPT_NOT_A_NODE} <= seen, seen
    assert any(x >= 0 for x in picks)
    eng.close()


def test_handmade_devices_priorities_and_reprieve(gpu_ok, oracle):
    def gpu_node(pods):
        return B.node({"cpu": 8000, "memory": 64 * GiB, "ephemeral-storage": 100 * GiB, "pods": 110, ("scalar", 0): 2, ("scalar", 1): 32 * GiB}, pods)

    def frac(mem, prio, idx, start=0, cpu=100):
        return B.pod([B.ctr({"cpu": cpu}, mem_limit=mem)], priority=prio, node_name=True, index=idx, start=start)

    def whole(prio, idx, start=0):
        return B.pod([B.ctr({"cpu": 100}, gpu_limit=1)], priority=prio, node_name=True, index=idx, start=start)

    nodes = [
        gpu_node([frac(10 * GiB, 10, 0, start=5), frac(4 * GiB, 20, 0, start=1), frac(12 * GiB, 10, 1, start=2)]),  # 0
        gpu_node([whole(10, 0), whole(200, 1)]),                                                                     # 1: only device 0 can be freed
        gpu_node([frac(8 * GiB, 300, 0), frac(8 * GiB, 300, 1), frac(8 * GiB, 300, 0), frac(8 * GiB, 300, 1)]),      # 2: nobody is less important
        gpu_node([frac(GiB, 10, 0, cpu=7900)]),                                                                      # 3: cpu is the problem
    ]
    pending = [
        B.pod([B.ctr({"cpu": 100}, gpu_limit=1)], priority=100),             # needs an idle device
        B.pod([B.ctr({"cpu": 100}, mem_limit=14 * GiB)], priority=100),      # needs 14 GiB on one device
        B.pod([B.ctr({"cpu": 4000})], priority=100),                         # no flex limit
        B.pod([B.ctr({"cpu": 100}, mem_limit=14 * GiB)], priority=5),        # less important than everybody
    ]
    cfg = fgs.config_plugins(2)
    snap, pend = B.build_snapshot(nodes), B.build_podset(pending)
    eng = _engine(snap.ptr, cfg)
    for p in range(len(pending)):
        st, vic, cnt = eng.select_victims(snap.ptr, pend.ptr, p)
        rst, rvic, rcnt = oracle.select_victims(snap.ptr, pend.ptr, p, cfg)
        assert np.array_equal(st, rst) and np.array_equal(vic, rvic) and np.array_equal(cnt, rcnt), p
        if p == 0:
            # a whole-GPU pod passes FlexGPU.Filter on the limit sums alone (flex_gpu.go:85-98), so every potential victim
            # of node 0 is reprieved again: "expected at least one victim pod" (preemption.go:578-580)
            assert int(st[0]) == L.FGS_PREEMPT_ERROR
        if p == 1:
            # 14 GiB on one 16 GiB device. Node 0: all three pods go, the priority-20 pod (4 GiB, device 0) comes back,
            # the 12 GiB pod of device 1 cannot (no device has 14 GiB left then), the 10 GiB pod of device 0 can (device 1
            # is empty now). Node 1: only the priority-10 whole-GPU pod is less important, and it has to go.
            assert st.tolist()[:3] == [0, 0, 1] and vic[:3].tolist() == [0, 0, 1] and vic[3:5].tolist() == [1, 0]
        if p == 3:
            assert (st == L.FGS_PREEMPT_NO_VICTIMS).all()
    eng.close()


def test_pick_one_node_criteria_on_the_gpu(gpu_ok, oracle):
    """every criterion of pickOneNodeForPreemption decides once: nodes that differ only in what that criterion looks at"""
    def n(pods):
        return B.node({"cpu": 1000, "memory": 64 * GiB, "pods": 110}, pods)

    def v(prio, start=0, cpu=600):
        return B.pod([B.ctr({"cpu": cpu})], priority=prio, node_name=True, start=start)

    pend = B.build_podset([B.pod([B.ctr({"cpu": 900})], priority=1000)])
    cfg = fgs.plugin_cfg(filter_plugins=("NodeResourcesFit",), score_plugins=(("NodeResourcesFit", 1),))
    INT32_MIN = -2147483648
    cases = [
        ([n([v(50)]), n([v(40)]), n([v(60)])], [1]),                                            # lowest highest-priority victim
        ([n([v(50, cpu=300), v(50, cpu=300)]), n([v(50, cpu=300), v(10, cpu=300)]), n([v(50, cpu=300), v(20, cpu=300)])], [1]),  # sum
        ([n([v(-1, cpu=300), v(-1, cpu=300)]), n([v(-1)])], [1]),                               # the +2^31 shift: fewer pods win
        ([n([v(7, cpu=300), v(INT32_MIN, cpu=300)]), n([v(7)])], [1]),                          # equal sums, fewer victims
        ([n([v(5, start=10)]), n([v(5, start=30)]), n([v(5, start=20)])], [1]),                 # latest start time
        ([n([v(5, start=9, cpu=300), v(5, start=30, cpu=300)]), n([v(5, start=10, cpu=300), v(5, start=11, cpu=300)])], [1]),  # EARLIEST of the top pods
        ([n([v(5, start=7)]), n([v(3)], ), n([v(5, start=7)]), n([v(3)])], [1, 3]),              # level to the end: a tie set
        ([n([v(2000)]), n([])], []),                                                            # nobody to evict
    ]
    for nodes, want in cases:
        snap = B.build_snapshot(nodes)
        eng = _engine(snap.ptr, cfg)
        st, vic, cnt, picked = eng.select_victims(snap.ptr, pend.ptr, 0, pick=True)
        rst, rvic, rcnt = oracle.select_victims(snap.ptr, pend.ptr, 0, cfg)
        assert np.array_equal(st, rst) and np.array_equal(vic, rvic)
        ties = np.nonzero(oracle.pick_preemption_node(snap.ptr, rst, rvic))[0].tolist()
        assert ties == want, (want, ties, st.tolist(), vic.tolist())
        assert picked == (want[0] if want else -1)
        eng.close()


def test_snapshot_pods_are_cached_until_the_node_table_changes(gpu_ok, oracle):
    """fgs_select_victims keeps the snapshot's pods on the device between calls; uploading the node table again (the
    snapshot object may be the same one, edited in place) makes it pack them afresh"""
    def n(pods):
        return B.node({"cpu": 1000, "memory": 64 * GiB, "pods": 110}, pods)

    def v(prio):
        return B.pod([B.ctr({"cpu": 600})], priority=prio, node_name=True)

    snap = B.build_snapshot([n([v(50)]), n([v(40)]), n([v(60)])])
    pend = B.build_podset([B.pod([B.ctr({"cpu": 900})], priority=1000), B.pod([B.ctr({"cpu": 900})], priority=45)])
    cfg = fgs.plugin_cfg(filter_plugins=("NodeResourcesFit",), score_plugins=(("NodeResourcesFit", 1),))
    eng = _engine(snap.ptr, cfg)
    for p, want in ((0, 1), (1, 1), (0, 1)):  # repeated calls, two preemptors: same cached columns
        st, vic, cnt, picked = eng.select_victims(snap.ptr, pend.ptr, p, pick=True)
        rst, rvic, rcnt = oracle.select_victims(snap.ptr, pend.ptr, p, cfg)
        assert np.array_equal(st, rst) and np.array_equal(vic, rvic) and picked == want
    # the same snapshot object, edited in place: node 2's pod becomes the least important one
    snap.ptr.contents.pods.priority[2] = 5
    eng.upload_nodes(fgs.Packer(2).nodes(snap.ptr))
    st, vic, cnt, picked = eng.select_victims(snap.ptr, pend.ptr, 0, pick=True)
    rst, rvic, rcnt = oracle.select_victims(snap.ptr, pend.ptr, 0, cfg)
    assert np.array_equal(st, rst) and np.array_equal(vic, rvic) and picked == 2
    eng.close()

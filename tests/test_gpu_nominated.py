"""Nominated pods: RunFilterPluginsWithNominatedPods (vendor/k8s.io/kubernetes/pkg/scheduler/framework/runtime/
framework.go:752-823) — a node that has pods nominated to it by an earlier preemption is filtered with those of
equal or higher priority added (addNominatedPods) and, if that passes, once more without them."""
import os

import numpy as np
import pytest

import builders as B
import flex_gpu_scheduler_b200 as fgs

pytestmark = pytest.mark.gpu
GiB = 1 << 30


def _engine(snapshot_ptr, cfg, force_int64=False):
    os.environ["FGS_FORCE_INT64"] = "1" if force_int64 else "0"
    try:
        eng = fgs.Engine(device=0, n_scalars=2)
    finally:
        os.environ.pop("FGS_FORCE_INT64", None)
    eng.set_plugins(cfg)
    eng.upload_nodes(fgs.Packer(2).nodes(snapshot_ptr))
    return eng


@pytest.mark.parametrize("force_int64", [False, True])
@pytest.mark.parametrize("cfg_id,with_uids", [(2, False), (2, True), (1, True)])
def test_synthetic_nominated_pods_match_the_oracle(gpu_ok, oracle, cfg_id, with_uids, force_int64):
    s = fgs.config_synth(cfg_id, n_nodes=400, n_pending=300, n_nominated=160, max_existing_pods=10)
    cfg = fgs.config_plugins(cfg_id)
    Q = int(s.nominated.contents.n_pods)
    rng = np.random.default_rng(7)
    nom_node = rng.integers(-1, 60, Q).astype(np.int32)  # several nominated pods per node, some without a node
    nom_uid = (1000 + np.arange(Q)).astype(np.int64)
    pend_uid = (5000 + np.arange(s.n_pending)).astype(np.int64)
    pend_uid[:40] = nom_uid[:40]  # some evaluated pods ARE nominated pods: they do not yield to themselves
    eng = _engine(s.snapshot, cfg, force_int64)
    nom_table = fgs.Packer(2).pods(s.nominated)
    eng.set_nominated(nom_table, nom_node, nom_uid if with_uids else None)
    table = fgs.Packer(2).pods(s.pending)
    if with_uids:
        fgs.Engine.with_uids(table, pend_uid)
    got = eng.evaluate(table, want_planes=True)
    ref = oracle.evaluate(s.snapshot, s.pending, cfg, nominated=(s.nominated, nom_node, nom_uid if with_uids else None,
                                                                pend_uid if with_uids else None))
    plain = oracle.evaluate(s.snapshot, s.pending, cfg)
    if cfg_id == 2:
        assert int((plain.status == 0).sum()) > int((ref.status == 0).sum()) + 30  # the nominated pods do take nodes away
    N = s.n_nodes
    assert np.array_equal(got.feasible(N), ref.status == 0)
    assert np.array_equal(got.feasible_count, ref.feasible_count)
    assert np.array_equal(got.pod_err, ref.pod_err)
    ok = (ref.pod_err == 0) & (ref.feasible_count > 0)
    assert np.array_equal(got.selected_score[ok], ref.max_total[ok])
    for p in np.nonzero(ok)[0]:
        n = int(got.selected_node[p])
        f = ref.status[p] == 0
        assert f[n] and int(ref.total[p][n]) == int(ref.max_total[p])
        assert np.array_equal(got.scores[p][:N][f], ref.total[p][f])
    # clearing the nominator restores the plain evaluation
    eng.set_nominated()
    again = eng.evaluate(table, want_planes=True)
    assert np.array_equal(again.feasible(N), plain.status == 0)
    eng.close()


def test_handmade_priorities_devices_and_pod_count(gpu_ok, oracle):
    def gpu_node(pods=None, max_pods=110):
        return B.node({"cpu": 8000, "memory": 64 * GiB, "ephemeral-storage": 100 * GiB, "pods": max_pods, ("scalar", 0): 2, ("scalar", 1): 32 * GiB},
                      pods or [])
    nodes = [gpu_node(), gpu_node(), gpu_node(max_pods=1), gpu_node()]
    nominated = [
        B.pod([B.ctr({"cpu": 6000})], priority=100),                          # node 0: leaves 2000m for pods of priority <= 100
        B.pod([B.ctr({"cpu": 100}, gpu_limit=1)], priority=50, index=0),     # node 1: holds device 0 …
        B.pod([B.ctr({"cpu": 100}, mem_limit=12 * GiB)], priority=50, index=1),  # … and 12 of device 1's 16 GiB
        B.pod([B.ctr({"cpu": 100})], priority=10),                            # node 2: the only pod slot
        B.pod([B.ctr({"cpu": 100}, mem_limit=GiB)], priority=10, index=7),    # node 3: an index NewGPUNode panics on
    ]
    nom_node = np.array([0, 1, 1, 2, 3], np.int32)
    pending = [
        B.pod([B.ctr({"cpu": 3000})], priority=100),   # yields to the 6000m pod: node 0 out
        B.pod([B.ctr({"cpu": 3000})], priority=101),   # outranks it: node 0 stays
        B.pod([B.ctr({"cpu": 100}, mem_limit=8 * GiB)], priority=50),   # node 1: device 0 taken, device 1 has 4 GiB → out
        B.pod([B.ctr({"cpu": 100}, mem_limit=4 * GiB)], priority=50),   # fits device 1
        B.pod([B.ctr({"cpu": 100}, gpu_limit=1)], priority=60),         # outranks both: node 1 stays
        B.pod([B.ctr({"cpu": 100})], priority=5),                       # node 2's slot is promised to a priority-10 pod
        B.pod([B.ctr({"cpu": 100}, mem_limit=GiB)], priority=5),        # node 3 with the bad index: the reference panics
    ]
    cfg = fgs.config_plugins(2)
    snap, pend, nom = B.build_snapshot(nodes), B.build_podset(pending), B.build_podset(nominated)
    eng = _engine(snap.ptr, cfg)
    eng.set_nominated(fgs.Packer(2).pods(nom.ptr), nom_node)
    got = eng.evaluate(fgs.Packer(2).pods(pend.ptr), want_planes=True)
    ref = oracle.evaluate(snap.ptr, pend.ptr, cfg, nominated=(nom.ptr, nom_node, None, None))
    feas = got.feasible(4)
    assert np.array_equal(feas, ref.status == 0)
    assert np.array_equal(got.pod_err, ref.pod_err)
    assert feas[0].tolist() == [False, True, True, True] and feas[1].tolist() == [True, True, True, True]
    assert not feas[2][1] and feas[3][1] and feas[4][1]
    assert not feas[5][2]
    assert int(got.pod_err[6]) & 2 and int(ref.pod_err[6]) & 2
    eng.close()

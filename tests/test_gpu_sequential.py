"""Sequential scheduling cycles on the GPU (fgs_schedule_sequential) against the oracle stepped pod by pod:
every cycle sees the node rows the previous commits left — assume/AddPod (framework/types.go:583-609) and the
device index FlexGPU.Reserve picked (flex_gpu.go:178-223)."""
import ctypes as C
import os

import numpy as np
import pytest

import builders as B
import seq_stepper as SS
import flex_gpu_scheduler_b200 as fgs

pytestmark = pytest.mark.gpu

GiB = 1 << 30


def _engine(nodes_or_snapshot, cfg, env=None, S=2):
    env = env or {}
    for k, v in env.items():
        os.environ[k] = v
    try:
        eng = fgs.Engine(device=0, n_scalars=S)
    finally:
        pass
    eng.set_plugins(cfg)
    pk = fgs.Packer(S)
    eng.upload_nodes(pk.nodes(nodes_or_snapshot))
    return eng


def _run(snapshot_ptr, pending_ptr, cfg, env=None, reserve_flex=True, quotas=None):
    env = env or {}
    for k, v in env.items():
        os.environ[k] = v
    try:
        eng = _engine(snapshot_ptr, cfg)
        table = fgs.Packer(2).pods(pending_ptr)
        res = eng.schedule_sequential(table, reserve_flex=reserve_flex, quotas=quotas)
    finally:
        for k in env:
            os.environ.pop(k, None)
    return eng, res


def _gpu_node(cpu=32000, mem=128 * GiB, gpus=2, gmem=32 * GiB, pods=None, max_pods=110):
    return B.node({"cpu": cpu, "memory": mem, "ephemeral-storage": 100 * GiB, "pods": max_pods, ("scalar", 0): gpus, ("scalar", 1): gmem},
                  pods or [])


def _frac(mem, cpu=500, ram=GiB):
    return B.pod([B.ctr({"cpu": cpu, "memory": ram}, mem_limit=mem)])


def _whole(cpu=1000, ram=2 * GiB):
    return B.pod([B.ctr({"cpu": cpu, "memory": ram}, gpu_limit=1)])


def test_handmade_queue_device_indexes_and_reserve_failure(gpu_ok, oracle):
    # two nodes x two 16 GiB devices
    nodes = [_gpu_node(), _gpu_node(cpu=16000)]
    queue = [
        _frac(10 * GiB),              # lands on some device 0
        _frac(10 * GiB),              # bin-pack: MemScore prefers the fuller node; 10+10 > 16 → device 1
        _frac(4 * GiB),               # first fit: device 0 has 6 GiB left
        _whole(),                     # passes the limit-sum filter everywhere; Reserve needs an idle device
        _whole(),
        _whole(),                     # by now at most ... whatever is left
        B.pod([B.ctr({"cpu": 100})]), # no flex limit: nothing to reserve
        B.pod([B.ctr({"cpu": 100}, gpu_limit=1, mem_limit=GiB)]),  # both limits: unresolvable everywhere
        _frac(17 * GiB),              # larger than any device
        _frac(6 * GiB), _frac(6 * GiB), _frac(6 * GiB), _frac(6 * GiB), _frac(6 * GiB), _frac(6 * GiB),
        _whole(), _whole(),
        B.pod([B.ctr({"cpu": 64000})]),  # no node has the cpu
    ]
    cfg = fgs.config_plugins(2)
    snap = B.build_snapshot(nodes)
    pend = B.build_podset(queue)
    eng, res = _run(snap.ptr, pend.ptr, cfg)
    final, stats = SS.follow(oracle, nodes, queue, cfg, res)
    assert stats["committed"] >= 8 and stats["unschedulable"] >= 3
    # the table the engine is left with equals a fresh upload of the stepped cluster: probe it with a batch
    probe = B.build_podset([_frac(2 * GiB), _frac(8 * GiB), _whole(), B.pod([B.ctr({"cpu": 100})])])
    got = eng.evaluate(fgs.Packer(2).pods(probe.ptr), want_planes=True)
    fin = B.build_snapshot(final)
    ref = oracle.evaluate(fin.ptr, probe.ptr, cfg)
    assert np.array_equal(got.feasible(len(final)), ref.status == 0)
    assert np.array_equal(got.feasible_count, ref.feasible_count)
    eng.close()


def test_whole_gpu_passes_filter_but_reserve_fails(gpu_ok, oracle):
    # one node, two devices each holding a fractional pod: the gpu limit sum (0 + 1 <= 2) lets a whole-GPU pod
    # through FlexGPU.Filter (flex_gpu.go:85-93) but GPUAssumeFitIndexes finds no idle device (gpu_node.go:163-177)
    held = [B.pod([B.ctr({"cpu": 100}, mem_limit=GiB)], node_name=True, index=0), B.pod([B.ctr({"cpu": 100}, mem_limit=GiB)], node_name=True, index=1)]
    nodes = [_gpu_node(pods=held)]
    queue = [_whole(), _frac(GiB)]
    cfg = fgs.config_plugins(2)
    snap, pend = B.build_snapshot(nodes), B.build_podset(queue)
    eng, res = _run(snap.ptr, pend.ptr, cfg)
    assert int(res.feasible_count[0]) == 1 and int(res.device_index[0]) == SS.RESERVE_FAIL and not res.committed[0]
    assert res.committed[1] and int(res.device_index[1]) == 0
    SS.follow(oracle, nodes, queue, cfg, res)
    eng.close()


# geometry variants: one thread-block cluster exchanging through distributed shared memory (the default for a
# cluster this small; seq_ctas is reported negative), one CTA, global mailboxes with the node state in shared
# memory, global mailboxes with the node state in global memory
# ... and the same with the commit applied after the verdict instead of being built ahead of it (FGS_SEQ_SPEC=0)
@pytest.mark.parametrize("env,ctas", [({}, -4), ({"FGS_SEQ_CTAS": "1"}, 1), ({"FGS_SEQ_CLUSTER": "0"}, 3),
                                      ({"FGS_SEQ_CTAS": "3", "FGS_SEQ_STATE_GLOBAL": "1"}, 3), ({"FGS_SEQ_SPEC": "0"}, -4),
                                      ({"FGS_SEQ_CTAS": "2"}, 2)])
@pytest.mark.parametrize("cfg_id", [1, 2])
def test_synthetic_cluster_follows_the_oracle(gpu_ok, oracle, cfg_id, env, ctas):
    s = fgs.config_synth(cfg_id, n_nodes=700, n_pending=260, max_existing_pods=6, pm_invalid_nodes=10)
    cfg = fgs.config_plugins(cfg_id)
    nodes = SS.snapshot_to_nodes(s.snapshot.contents)
    queue = SS.podset_to_pods(s.pending.contents)
    # the converter reproduces the generator's NodeInfo aggregates
    rebuilt = B.build_snapshot(nodes)
    N, S = s.n_nodes, 2
    assert np.array_equal(np.ctypeslib.as_array(rebuilt.struct.requested, shape=(N * (3 + S),)),
                          np.ctypeslib.as_array(s.snapshot.contents.requested, shape=(N * (3 + S),)))
    assert np.array_equal(np.ctypeslib.as_array(rebuilt.struct.nonzero, shape=(N * 2,)),
                          np.ctypeslib.as_array(s.snapshot.contents.nonzero, shape=(N * 2,)))
    eng, res = _run(s.snapshot, s.pending, cfg, env)
    t = eng.timings()
    assert t["seq_ctas"] == ctas
    final, stats = SS.follow(oracle, nodes, queue, cfg, res)
    assert stats["committed"] > 100
    # batch evaluation of a probe against the table the commits left behind
    probe = fgs.config_synth(cfg_id, n_nodes=8, n_pending=64, seed_pods=77)
    got = eng.evaluate(fgs.Packer(2).pods(probe.pending), want_planes=True)
    fin = B.build_snapshot(final)
    ref = oracle.evaluate(fin.ptr, probe.pending, cfg)
    assert np.array_equal(got.feasible(N), ref.status == 0)
    ok = (ref.pod_err == 0) & (ref.feasible_count > 0)
    assert np.array_equal(got.selected_score[ok], ref.max_total[ok])
    eng.close()


def test_cycles_that_end_in_an_error_commit_nothing(gpu_ok, oracle):
    # nodes on which NewGPUNode panics (no gpus, a bad index annotation) put most cycles into the error state:
    # FlexGPU.Filter reaches them for flex pods, FlexGPU.Score for every pod kind once they are feasible
    s = fgs.config_synth(2, n_nodes=300, n_pending=200, max_existing_pods=4, pm_panic_nodes=4, pm_nongpu_nodes=30, pm_invalid_nodes=10)
    cfg = fgs.config_plugins(2)
    nodes = SS.snapshot_to_nodes(s.snapshot.contents)
    queue = SS.podset_to_pods(s.pending.contents)
    eng, res = _run(s.snapshot, s.pending, cfg)
    final, stats = SS.follow(oracle, nodes, queue, cfg, res)
    assert stats["errors"] > 20
    assert not res.committed[res.pod_err != 0].any()
    eng.close()


def test_most_allocated_and_allocatable_scores_without_reserve_plugin(gpu_ok, oracle):
    s = fgs.config_synth(2, n_nodes=300, n_pending=150, max_existing_pods=4)
    cfg = fgs.plugin_cfg(filter_plugins=("FlexGPU", "NodeResourcesFit"),
                         score_plugins=(("FlexGPU", 2), ("NodeResourcesFit", 3), ("NodeResourcesAllocatable", 1)), fit_strategy="MostAllocated",
                         fit_resources=(("cpu", 2), ("memory", 1)), alloc_mode="Most")
    nodes = SS.snapshot_to_nodes(s.snapshot.contents)
    queue = SS.podset_to_pods(s.pending.contents)
    eng, res = _run(s.snapshot, s.pending, cfg, reserve_flex=False)
    assert (res.device_index == SS.RESERVE_SKIPPED).all()
    final, stats = SS.follow(oracle, nodes, queue, cfg, res, reserve_flex=False)
    assert stats["committed"] > 50
    eng.close()


@pytest.mark.parametrize("env", [{}, {"FGS_SEQ_CLUSTER": "0"}, {"FGS_SEQ_CTAS": "1"}, {"FGS_SEQ_SPEC": "0"}])
def test_elastic_quota_prefilter_and_reserve_between_cycles(gpu_ok, oracle, env):
    # CapacityScheduling in the profile: every cycle starts with its PreFilter against the quotas as the earlier
    # commits left them; a committed pod is added to its namespace's Used (capacity_scheduling.go:201-275,340-353)
    s = fgs.config_synth(4, n_nodes=600, n_pending=400, n_namespaces=12, max_existing_pods=4)
    cfg = fgs.config_plugins(4)
    nodes = SS.snapshot_to_nodes(s.snapshot.contents)
    queue = SS.podset_to_pods(s.pending.contents)
    qstate = SS.QuotaState(s.quotas.contents)
    # tighten Max so that namespaces fill up during the queue, and Σmin so that the aggregate check bites late
    qstate.max[:, :2] = qstate.used[:, :2] + np.array([[40000, 400 * GiB]])
    qstate.min[:, 1] = qstate.used[:, 1] + 100 * GiB
    qs = qstate.struct()
    eng, res = _run(s.snapshot, s.pending, cfg, env, quotas=C.pointer(qs))
    final, stats = SS.follow(oracle, nodes, queue, cfg, res, quotas=qstate)
    assert stats["committed"] > 50 and stats["quota_rejected"] > 20, stats
    assert {0, 1, 2} <= set(int(x) for x in res.quota_code)
    assert np.array_equal(res.quota_used, qstate.used)
    assert np.array_equal(res.quota_used_present, qstate.present)
    eng.close()


def test_reserve_batch_matches_oracle(gpu_ok, oracle):
    s = fgs.config_synth(2, n_nodes=400, n_pending=500, pm_panic_nodes=30, pm_nongpu_nodes=50)
    eng = _engine(s.snapshot, fgs.config_plugins(2))
    rng = np.random.default_rng(5)
    nodes = rng.integers(0, s.n_nodes, s.n_pending).astype(np.int32)
    nodes[::17] = -1
    dev = eng.reserve(fgs.Packer(2).pods(s.pending), nodes)
    for p in range(s.n_pending):
        want = SS.RESERVE_SKIPPED if nodes[p] < 0 else oracle.reserve_index(s.snapshot, s.pending, p, int(nodes[p]))
        assert int(dev[p]) == want, (p, int(nodes[p]), int(dev[p]), want)
    assert {SS.RESERVE_FAIL, SS.RESERVE_NONE, SS.RESERVE_CONFLICT, SS.RESERVE_PANIC} <= set(int(x) for x in dev) and (dev >= 0).any()
    eng.close()


def test_sequential_without_commits_equals_the_batch(gpu_ok):
    # a queue in which nothing can be placed twice... simpler: pods that fit nowhere leave the table alone, and the
    # first cycle of any queue equals the batched evaluation of its first pod
    s = fgs.config_synth(2, n_nodes=1500, n_pending=40)
    cfg = fgs.config_plugins(2)
    eng, res = _run(s.snapshot, s.pending, cfg)
    eng2 = _engine(s.snapshot, cfg)
    batch = eng2.evaluate(fgs.Packer(2).pods(s.pending))
    assert int(res.feasible_count[0]) == int(batch.feasible_count[0])
    assert int(res.selected_node[0]) == int(batch.selected_node[0])  # same tie-break hash
    assert int(res.selected_score[0]) == int(batch.selected_score[0])
    eng.close()
    eng2.close()

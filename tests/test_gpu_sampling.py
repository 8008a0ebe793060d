"""Node sampling (fgs_eval_opts.sample): the reference's findNodesThatPassFilters walk at one worker
(vendor/k8s.io/kubernetes/pkg/scheduler/generic_scheduler.go:171-191,266-338) against the oracle's restatement —
the kept feasible set, the nodes each walk counted, nextStartNodeIndex carried from pod to pod, and
NormalizeScore / selectHost over the kept nodes only."""
import os

import numpy as np
import pytest

import flex_gpu_scheduler_b200 as fgs

pytestmark = pytest.mark.gpu


def _engine(s, cfg, force_int64):
    os.environ["FGS_FORCE_INT64"] = "1" if force_int64 else "0"
    try:
        eng = fgs.Engine(device=0, n_scalars=2)
    finally:
        os.environ.pop("FGS_FORCE_INT64", None)
    eng.set_plugins(cfg)
    eng.upload_nodes(fgs.Packer(2).nodes(s.snapshot))
    return eng


@pytest.mark.parametrize("force_int64", [False, True])
@pytest.mark.parametrize("cfg_id,n_nodes,n_pods,pct,start", [(2, 1500, 300, 0, 0), (2, 1500, 200, 30, 1234), (1, 2100, 150, 0, 2099),
                                                           (2, 90, 64, 0, 5), (2, 700, 120, 100, 3)])
def test_sampled_evaluation_matches_the_oracle(gpu_ok, oracle, cfg_id, n_nodes, n_pods, pct, start, force_int64):
    s = fgs.config_synth(cfg_id, n_nodes=n_nodes, n_pending=n_pods, pm_invalid_nodes=5)
    cfg = fgs.config_plugins(cfg_id)
    eng = _engine(s, cfg, force_int64)
    got = eng.evaluate(fgs.Packer(2).pods(s.pending), want_planes=True, sample=1, percentage_of_nodes_to_score=pct, start_node_index=start)
    t = eng.timings()
    assert t["reduce_lookup"] and t["emit_int32"] == (not force_int64)
    ref = oracle.evaluate(s.snapshot, s.pending, cfg, sample=(pct, start))
    K = oracle.num_feasible_nodes_to_find(n_nodes, pct)
    assert np.array_equal(got.feasible(n_nodes), ref.status == 0)
    assert np.array_equal(got.feasible_count, ref.feasible_count)
    assert int(got.feasible_count.max()) <= K
    assert np.array_equal(got.processed_nodes, ref.processed)
    assert got.next_start_node_index == ref.next_start
    assert np.array_equal(got.pod_err, ref.pod_err)
    ok = (ref.pod_err == 0) & (ref.feasible_count > 0)
    assert np.array_equal(got.selected_score[ok], ref.max_total[ok])
    for p in np.nonzero(ok)[0]:
        n = int(got.selected_node[p])
        assert ref.status[p][n] == 0 and int(ref.total[p][n]) == int(ref.max_total[p])
        f = ref.status[p] == 0
        assert np.array_equal(got.scores[p][:n_nodes][f], ref.total[p][f])
    if pct < 100 and n_nodes >= 100:
        assert (got.processed_nodes < n_nodes).any()  # the walk really stopped early for some pods
    eng.close()


def test_sampling_is_a_chain_over_batches(gpu_ok, oracle):
    # two batches evaluated one after the other continue the same walk as one batch of both
    s = fgs.config_synth(2, n_nodes=1200, n_pending=160)
    cfg = fgs.config_plugins(2)
    eng = _engine(s, cfg, False)
    whole = eng.evaluate(fgs.Packer(2).pods(s.pending), sample=1, start_node_index=17)
    a = eng.evaluate(fgs.Packer(2).pods(s.pending, 0, 90), sample=1, start_node_index=17)
    b = eng.evaluate(fgs.Packer(2).pods(s.pending, 90, 160), sample=1, start_node_index=a.next_start_node_index)
    assert b.next_start_node_index == whole.next_start_node_index
    assert np.array_equal(np.concatenate([a.processed_nodes, b.processed_nodes]), whole.processed_nodes)
    # (the tie-break hash is seeded with the pod's index in its batch: compare scores, not the tie choice)
    assert np.array_equal(np.concatenate([a.selected_score, b.selected_score]), whole.selected_score)
    assert np.array_equal(np.concatenate([a.feasible_count, b.feasible_count]), whole.feasible_count)
    # without sampling nothing of this is touched
    full = eng.evaluate(fgs.Packer(2).pods(s.pending))
    assert full.next_start_node_index == -1 and (full.processed_nodes == 1200).all()
    assert (full.feasible_count >= whole.feasible_count).all()
    eng.close()


@pytest.mark.parametrize("cfg_id,n_nodes,n_pods,pct,start,kw", [
    (2, 1500, 5000, 0, 77, {}),
    (2, 5000, 3000, 10, 4999, {}),
    (1, 2100, 2500, 0, 0, {}),
    (2, 3000, 4100, 5, 1, dict(pm_nongpu_nodes=700, pm_quirk_nodes=100)),   # many pods with fewer than K feasible nodes: their walk leaves the index alone
    (2, 16384, 1100, 3, 16000, {}),                                         # the largest table whose merge arrays fit in shared memory
    (2, 21000, 1300, 2, 20999, {}),                                         # beyond: they live in global scratch
])
def test_long_batches_walk_in_blocks_and_land_where_the_chain_does(gpu_ok, oracle, cfg_id, n_nodes, n_pods, pct, start, kw):
    """1 024 pods or more on up to 65 528 nodes: the walk is composed per block of 256 pods (sample_compose), the blocks are hopped
    over (sample_block_starts) and walked at once (sample_chain). Same kept sets, processed counts and final index as the
    oracle's one-pod-after-the-other walk, and as the serial kernel (FGS_SAMPLE_SERIAL=1)."""
    s = fgs.config_synth(cfg_id, n_nodes=n_nodes, n_pending=n_pods, pm_invalid_nodes=5, **kw)
    cfg = fgs.config_plugins(cfg_id)
    ref = oracle.evaluate(s.snapshot, s.pending, cfg, sample=(pct, start))
    got = {}
    for serial in (False, True):
        if serial:
            os.environ["FGS_SAMPLE_SERIAL"] = "1"
        try:
            eng = _engine(s, cfg, False)
            l0 = eng.timings()["kernel_launches"]
            r = eng.evaluate(fgs.Packer(2).pods(s.pending), want_planes=True, sample=1, percentage_of_nodes_to_score=pct, start_node_index=start)
            got[serial] = (r, eng.timings()["kernel_launches"] - l0)
            eng.close()
        finally:
            os.environ.pop("FGS_SAMPLE_SERIAL", None)
    assert got[False][1] == got[True][1] + 2  # sample_compose and sample_block_starts ran
    for serial in (False, True):
        r = got[serial][0]
        assert np.array_equal(r.processed_nodes, ref.processed), serial
        assert r.next_start_node_index == ref.next_start, serial
        assert np.array_equal(r.feasible(n_nodes), ref.status == 0), serial
        assert np.array_equal(r.feasible_count, ref.feasible_count) and np.array_equal(r.pod_err, ref.pod_err)
        ok = (ref.pod_err == 0) & (ref.feasible_count > 0)
        assert np.array_equal(r.selected_score[ok], ref.max_total[ok])
    assert np.array_equal(got[False][0].selected_node, got[True][0].selected_node)
    assert (ref.processed < n_nodes).any() and (ref.processed == n_nodes).any() or n_nodes > 10000 or cfg_id == 1 or not kw

"""fgs_score_masked: Score / NormalizeScore / selectHost under the framework's own feasible list (the nodes that passed
EVERY filter plugin of the profile, vendor/k8s.io/kubernetes/pkg/scheduler/generic_scheduler.go:420-426) against the oracle
given the same list."""
import os

import numpy as np
import pytest

import flex_gpu_scheduler_b200 as fgs
from flex_gpu_scheduler_b200 import _lib as L

pytestmark = pytest.mark.gpu


def _engine(s, cfg, force_int64, filt):
    os.environ["FGS_FORCE_INT64"] = "1" if force_int64 else "0"
    os.environ["FGS_FILTER"] = filt
    try:
        eng = fgs.Engine(device=0, n_scalars=2, total_nodes=s.n_nodes)
    finally:
        os.environ.pop("FGS_FORCE_INT64", None)
        os.environ.pop("FGS_FILTER", None)
    eng.set_plugins(cfg)
    eng.upload_nodes(fgs.Packer(2).nodes(s.snapshot))
    return eng


def _check(got, ref, N):
    feas = ref.status == 0
    assert np.array_equal(got.feasible(N), feas)
    assert np.array_equal(got.feasible_count, ref.feasible_count)
    assert np.array_equal(got.pod_err, ref.pod_err)
    ok = (ref.pod_err == 0) & (ref.feasible_count >= 1)
    m = feas & ok[:, None]
    assert np.array_equal(got.scores[:, :N][m], ref.total[m])
    idx = np.nonzero(ok)[0]
    sel = got.selected_node
    assert np.all(sel[ref.feasible_count == 0] == -1)
    assert np.all(feas[idx, sel[idx]])
    assert np.array_equal(ref.total[idx, sel[idx]], ref.max_total[idx])
    assert np.array_equal(got.selected_score[idx], ref.max_total[idx])


CASES = [
    ("c2", dict(cfg=2, n_nodes=1300, n_pending=400, pm_quirk_nodes=100, pm_panic_nodes=30), None),
    ("c1", dict(cfg=1, n_nodes=1100, n_pending=300), None),
    ("all-three", dict(cfg=2, n_nodes=700, n_pending=257),
     dict(score_plugins=(("FlexGPU", 2), ("NodeResourcesFit", 1), ("NodeResourcesAllocatable", 5)), alloc_mode="Most")),
]


@pytest.mark.parametrize("force_int64,filt", [(False, "lookup"), (False, "compare"), (True, "compare")])
@pytest.mark.parametrize("name,skw,pkw", CASES, ids=[c[0] for c in CASES])
def test_score_masked_matches_the_oracle_given_the_same_list(gpu_ok, oracle, name, skw, pkw, force_int64, filt):
    skw = dict(skw)
    cfg_id = skw.pop("cfg")
    s = fgs.config_synth(cfg_id, **skw)
    cfg = fgs.config_plugins(cfg_id) if pkw is None else fgs.plugin_cfg(**pkw)
    N, P = s.n_nodes, s.n_pending
    eng = _engine(s, cfg, force_int64, filt)
    table = fgs.Packer(2).pods(s.pending)
    rng = np.random.default_rng(7 + N)
    keep = rng.uniform(0.3, 0.9, size=(P, 1))  # each pod keeps 30-90 % of the nodes
    mask = rng.random((P, N)) < keep
    mask[3] = False        # another plugin rejected every node
    mask[5] = True         # … or none at all
    flt = eng.evaluate(table, want_planes=True, skip_pass_b=1)  # the Filter half of the cycle
    full = oracle.evaluate(s.snapshot, s.pending, cfg, workers=4)
    assert np.array_equal(flt.feasible(N), full.status == 0)
    got = eng.score_masked(mask, want_planes=True)
    ref = oracle.evaluate(s.snapshot, s.pending, cfg, workers=4, ext_mask=mask.astype(np.uint8))
    assert (ref.feasible_count < full.feasible_count).any()
    assert (ref.max_total[ref.feasible_count > 1] != full.max_total[ref.feasible_count > 1]).any()  # the normalisers really moved
    _check(got, ref, N)
    # the list can only shrink: a second, narrower mask on the same batch
    mask2 = mask & (rng.random((P, N)) < 0.7)
    got2 = eng.score_masked(mask2, want_planes=True)
    ref2 = oracle.evaluate(s.snapshot, s.pending, cfg, workers=4, ext_mask=mask2.astype(np.uint8))
    _check(got2, ref2, N)
    # a new evaluation of the same table is unmasked again
    again = eng.evaluate(table, want_planes=True)
    _check(again, full, N)
    eng.close()


def test_score_masked_needs_a_filtered_batch(gpu_ok):
    s = fgs.config_synth(2, n_nodes=200, n_pending=64)
    eng = fgs.Engine(device=0)
    eng.set_plugins(fgs.config_plugins(2))
    eng.upload_nodes(fgs.Packer(2).nodes(s.snapshot))
    eng.n_pods = 64
    with pytest.raises(fgs.FgsError):
        eng.score_masked(np.ones((64, 200), bool))
    eng.close()

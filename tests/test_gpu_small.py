"""Small batches — the reference's own call shape is one pod per scheduling cycle
(vendor/k8s.io/kubernetes/pkg/scheduler/scheduler.go:425-455): fgs_evaluate with at most 256 pods runs the whole cycle of
every pod in ONE launch (a CTA per pod). Its answers against the oracle, and against the general path bit for bit —
selectHost's choice included, the two build the same key."""
import os

import numpy as np
import pytest

import flex_gpu_scheduler_b200 as fgs
from flex_gpu_scheduler_b200 import _lib as L

pytestmark = pytest.mark.gpu

CASES = [
    ("c2", dict(cfg=2, n_nodes=5000, pm_quirk_nodes=60), None),
    ("c2-panics-flex-first", dict(cfg=2, n_nodes=900, pm_quirk_nodes=150, pm_panic_nodes=80, pm_nongpu_nodes=60, pm_invalid_nodes=20),
     dict(filter_plugins=("FlexGPU", "NodeResourcesFit"))),
    ("c1-allocatable", dict(cfg=1, n_nodes=1300), None),
    ("most-allocated-weights", dict(cfg=2, n_nodes=700),
     dict(score_plugins=(("FlexGPU", 3), ("NodeResourcesFit", 2)), fit_strategy="MostAllocated", fit_resources=(("cpu", 3), ("memory", 7)))),
    ("all-three", dict(cfg=2, n_nodes=1100),
     dict(score_plugins=(("FlexGPU", 2), ("NodeResourcesFit", 1), ("NodeResourcesAllocatable", 5)), alloc_mode="Most")),
    ("fit-only", dict(cfg=2, n_nodes=600), dict(filter_plugins=("NodeResourcesFit",), score_plugins=(("NodeResourcesFit", 1),))),
    # a long node list: up to eight pods take a cluster of CTAs each (two 1024-node units per CTA here, merged through DSMEM)
    ("c1-long-node-list", dict(cfg=1, n_nodes=15400, pm_quirk_nodes=40, pm_panic_nodes=20), None),
]


def _engine(s, cfg, small, direct=None):
    os.environ["FGS_SMALL"] = "1" if small else "0"
    if direct is not None:
        os.environ["FGS_SMALL_DIRECT"] = "1" if direct else "0"
    try:
        eng = fgs.Engine(device=0, n_scalars=2, total_nodes=s.n_nodes)
    finally:
        os.environ.pop("FGS_SMALL", None)
        os.environ.pop("FGS_SMALL_DIRECT", None)
    eng.set_plugins(cfg)
    eng.upload_nodes(fgs.Packer(2).nodes(s.snapshot))
    return eng


@pytest.mark.parametrize("direct", [False, True], ids=["level-tables", "compare"])
@pytest.mark.parametrize("name,skw,pkw", CASES, ids=[c[0] for c in CASES])
def test_one_launch_cycles_match_the_oracle_and_the_general_path(gpu_ok, oracle, name, skw, pkw, direct):
    """direct: the kernel's Filter without level tables (what it does when the node table changed since they were built)"""
    skw = dict(skw)
    cfg_id = skw.pop("cfg")
    s = fgs.config_synth(cfg_id, n_pending=400, **skw)
    cfg = fgs.config_plugins(cfg_id) if pkw is None else fgs.plugin_cfg(**pkw)
    N = s.n_nodes
    small, general = _engine(s, cfg, True, direct), _engine(s, cfg, False)
    ref_all = oracle.evaluate(s.snapshot, s.pending, cfg, workers=4)
    seen_err = 0
    for p0, P in ((0, 1), (1, 1), (7, 2), (10, 63), (100, 64), (40, 17), (0, 256), (130, 255)):
        t = fgs.Packer(2).pods(s.pending, p0, p0 + P)
        a = small.evaluate(t, want_planes=True)
        assert small.timings()["small"] and small.timings()["reduce_lookup"] == (not direct)
        b = general.evaluate(t, want_planes=True)
        assert not general.timings()["small"]
        ref_status = ref_all.status[p0:p0 + P]
        feas = ref_status == 0
        assert np.array_equal(a.feasible(N), feas)
        assert np.array_equal(a.feasible_count, ref_all.feasible_count[p0:p0 + P])
        assert np.array_equal(a.pod_err, ref_all.pod_err[p0:p0 + P])
        seen_err |= int(a.pod_err.max())
        ok = (a.pod_err == 0) & (a.feasible_count >= 1)
        m = feas & ok[:, None]
        assert np.array_equal(a.scores[:, :N][m], ref_all.total[p0:p0 + P][m])
        idx = np.nonzero(ok)[0]
        assert np.array_equal(a.selected_score[idx], ref_all.max_total[p0:p0 + P][idx])
        assert np.all(a.selected_node[a.feasible_count == 0] == -1)
        # … and the general path, including which node of the tie set is chosen
        for k in ("feasible_count", "selected_node", "selected_score", "pod_err"):
            assert np.array_equal(getattr(a, k), getattr(b, k)), k
        assert np.array_equal(a.feasible(N), b.feasible(N))
        assert np.array_equal(a.scores[:, :N], b.scores[:, :N])
        # the staged batch is as the general path leaves it: the status plane, a per-plugin emit pass, masked scoring
        assert np.array_equal(small.status_plane(0), ref_status)
    if "panics" in name:
        assert seen_err & L.FGS_PODERR_REF_PANIC
    t = fgs.Packer(2).pods(s.pending, 0, 32)
    a = small.evaluate(t, want_planes=True)
    assert small.timings()["small"]
    if cfg.score_mask & L.FGS_PL_FLEXGPU:
        one = small.evaluate(t, want_planes=True, emit_mask=L.FGS_PL_FLEXGPU)  # options: the general path takes over
        assert not small.timings()["small"]
        two = general.evaluate(t, want_planes=True, emit_mask=L.FGS_PL_FLEXGPU)
        assert np.array_equal(one.scores[:, :N], two.scores[:, :N])
    a = small.evaluate(t, want_planes=True)
    rng = np.random.default_rng(5)
    mask = rng.random((32, N)) < 0.6
    got = small.score_masked(mask, want_planes=True)
    ref = oracle.evaluate(s.snapshot, s.pending, cfg, 0, 32, workers=4, ext_mask=mask.astype(np.uint8))
    assert np.array_equal(got.feasible(N), ref.status == 0) and np.array_equal(got.feasible_count, ref.feasible_count)
    okm = (ref.pod_err == 0) & (ref.feasible_count >= 1)
    assert np.array_equal(got.selected_score[okm], ref.max_total[okm])
    big = small.evaluate(fgs.Packer(2).pods(s.pending, 0, 257))
    assert not small.timings()["small"]
    assert np.array_equal(big.feasible_count, ref_all.feasible_count[:257])
    small.close()
    general.close()


def test_cycles_between_node_changes_need_no_table_rebuild(gpu_ok, oracle):
    """The reference's loop: schedule one pod, bind it (the node's NodeInfo changes), schedule the next
    (scheduler.go:425-455, cache.go:197-276). After fgs_sync_nodes the level tables describe the OLD node table; a small
    batch is answered by the compare form of the one-launch kernel instead of rebuilding them, a large one rebuilds."""
    import ctypes as C
    S, N = 2, 2500
    s = fgs.config_synth(2, n_nodes=N, n_pending=300, pm_quirk_nodes=40)
    cfg = fgs.config_plugins(2)
    # a private copy of the snapshot whose aggregates this test edits the way a bind does (NodeInfo.AddPod:
    # Requested / NonZeroRequested grow, Generation moves)
    snap = L.Snapshot()
    C.memmove(C.byref(snap), s.snapshot, C.sizeof(L.Snapshot))
    req = np.ctypeslib.as_array(snap.requested, shape=(N, 3 + S)).copy()
    nz = np.ctypeslib.as_array(snap.nonzero, shape=(N, 2)).copy()
    gen = np.arange(1, N + 1, dtype=np.int64)
    snap.requested, snap.nonzero, snap.generation = (a.ctypes.data_as(L.p_i64) for a in (req, nz, gen))
    snap_p = C.pointer(snap)
    eng = fgs.Engine(device=0)
    eng.set_plugins(cfg)
    pk = fgs.Packer(S)

    def check(p0, p1, want_lookup):
        eng.sync_nodes(pk, snap_p)
        got = eng.evaluate(fgs.Packer(S).pods(s.pending, p0, p1), want_planes=True)
        tm = eng.timings()
        assert tm["small"] and tm["reduce_lookup"] == want_lookup, tm
        ref = oracle.evaluate(snap_p, s.pending, cfg, p0, p1, workers=2)
        assert np.array_equal(got.feasible(N), ref.status == 0)
        assert np.array_equal(got.feasible_count, ref.feasible_count) and np.array_equal(got.pod_err, ref.pod_err)
        ok = (ref.pod_err == 0) & (ref.feasible_count >= 1)
        m = (ref.status == 0) & ok[:, None]
        assert np.array_equal(got.scores[:, :N][m], ref.total[m])
        assert np.array_equal(got.selected_score[ok], ref.max_total[ok])
        idx = np.nonzero(ok)[0]
        assert np.array_equal(ref.total[idx, got.selected_node[idx]], ref.max_total[idx])
        return got

    check(0, 1, False)        # nothing built yet: one pod does not pay for the tables
    check(0, 256, True)       # 256 pods do (256 x 2500 is still one launch)
    check(5, 25, True)        # ... and now they are current
    check(5, 6, False)        # (up to eight pods on 2048+ nodes: a cluster of CTAs per pod, which never needs them)
    g = N + 1
    rng = np.random.default_rng(4242)
    for cycle in range(12):   # the pod of the last cycle is bound to the node chosen for it before the next one runs
        got = check(cycle, cycle + 1 + (cycle % 3) * 7, False)
        i = int(got.selected_node[0]) if got.feasible_count[0] > 0 else int(rng.integers(N))
        req[i, 0] += 500
        req[i, 1] += 256 << 20
        nz[i, 0] += 500
        nz[i, 1] += 256 << 20
        gen[i] = g
        g += 1
    check(0, 256, True)
    eng.close()


def test_level_tables_built_by_either_sort_answer_alike(gpu_ok, oracle):
    """the level tables of a node table of up to 16 384 rows are sorted in shared memory in one launch, larger ones by the
    library's radix sort; both orders are (key, node index), so every answer — tie-breaks included — is the same"""
    s = fgs.config_synth(2, n_nodes=9000, n_pending=2000, pm_quirk_nodes=150, pm_panic_nodes=40, pm_nongpu_nodes=60, pm_invalid_nodes=20)
    cfg = fgs.config_plugins(2)
    got = []
    for mode in (None, "cub"):
        if mode:
            os.environ["FGS_FT_SORT"] = mode
        try:
            eng = fgs.Engine(device=0)
            eng.set_plugins(cfg)
            eng.upload_nodes(fgs.Packer(2).nodes(s.snapshot))
            got.append(eng.evaluate(fgs.Packer(2).pods(s.pending), want_planes=True))
            assert eng.timings()["reduce_lookup"]
            eng.close()
        finally:
            os.environ.pop("FGS_FT_SORT", None)
    a, b = got
    N = s.n_nodes
    for k in ("feasible_count", "selected_node", "selected_score", "pod_err"):
        assert np.array_equal(getattr(a, k), getattr(b, k)), k
    assert np.array_equal(a.feasible(N), b.feasible(N)) and np.array_equal(a.scores[:, :N], b.scores[:, :N])
    ref = oracle.evaluate(s.snapshot, s.pending, cfg, workers=4)
    assert np.array_equal(a.feasible(N), ref.status == 0) and np.array_equal(a.feasible_count, ref.feasible_count)

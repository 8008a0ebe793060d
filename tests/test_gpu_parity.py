"""GPU parity tests: the CUDA path, called through the libfgs C ABI, against the CPU oracle on the same
seeded object snapshots, and against the golden vectors transcribed from the reference's unit tests.
Integer work: every comparison is bit-exact."""
import os

import numpy as np
import pytest

import builders as B
import golden_cases as G
import flex_gpu_scheduler_b200 as fgs
from flex_gpu_scheduler_b200 import _lib as L

pytestmark = pytest.mark.gpu

FLEX, FIT, ALLOC = L.FGS_PL_FLEXGPU, L.FGS_PL_FIT, L.FGS_PL_ALLOCATABLE


def run_engine(synth_or_pair, cfg, S=2, want_planes=True, **opts):
    snapshot, pending = synth_or_pair
    eng = fgs.Engine(device=0, n_scalars=S, total_nodes=int(snapshot.contents.n_nodes))
    eng.set_plugins(cfg)
    pk = fgs.Packer(S)
    eng.upload_nodes(pk.nodes(snapshot))
    res = eng.evaluate(pk.pods(pending), want_planes=want_planes, **opts)
    return eng, pk, res


PATHS_TAKEN = {}  # test id → (reduce_int32, emit_int32) of the unforced run


def check_against_oracle(oracle, snapshot, pending, cfg, S=2, per_plugin=True, tag=None):
    """Both kernel families must agree with the oracle: the default run (int32 specialisation when the batch
    qualifies) and a run with FGS_FORCE_INT64=1 (always the int64 kernels)."""
    ref = oracle.evaluate(snapshot, pending, cfg, workers=4, per_plugin=per_plugin)
    res = None
    # … and both forms of pass A: Filter as table look-ups (the default) and the compare kernels
    for filt in ("lookup", "compare"):
        for force in ("0", "1"):
            os.environ["FGS_FORCE_INT64"] = force
            os.environ["FGS_FILTER"] = filt
            try:
                res_i, took, looked_up = _check_one(ref, snapshot, pending, cfg, S, per_plugin)
            finally:
                os.environ.pop("FGS_FORCE_INT64", None)
                os.environ.pop("FGS_FILTER", None)
            assert looked_up == (filt == "lookup")
            if filt == "lookup":
                assert not took[0]
                if force == "0":
                    res = res_i
                    if tag:
                        PATHS_TAKEN[tag + "/lookup"] = took
            elif force == "0":
                if tag:
                    PATHS_TAKEN[tag] = took
            else:
                assert took == (False, False)
    return ref, res


def _check_one(ref, snapshot, pending, cfg, S, per_plugin):
    N = int(snapshot.contents.n_nodes)
    eng, pk, res = run_engine((snapshot, pending), cfg, S=S)
    t = eng.timings()
    took = (t["reduce_int32"], t["emit_int32"])
    looked_up = t["reduce_lookup"]
    feas_ref = ref.status == 0
    # identical feasible set, counts and error flags
    feas = res.feasible(N)
    assert np.array_equal(feas, feas_ref)
    assert np.array_equal(res.feasible_count, ref.feasible_count)
    assert np.array_equal(res.pod_err, ref.pod_err)
    # identical Filter status plane (merged, with the failing plugin) and per-plugin verdicts
    st = eng.status_plane(0)
    assert np.array_equal(st, ref.status)
    ok = ref.pod_err == 0  # scores of pods whose cycle errors out in the reference are not defined
    scored = ok & (ref.feasible_count >= 1)
    tot = res.scores[:, :N]
    m = feas_ref & scored[:, None]
    assert np.array_equal(tot[m], ref.total[m])
    # selectHost: chosen node lies in the reference's max-score tie set
    has = ok & (ref.feasible_count >= 1)
    sel = res.selected_node
    assert np.all(sel[ref.feasible_count == 0] == -1)
    idx = np.nonzero(has)[0]
    assert np.all(sel[idx] >= 0)
    assert np.all(feas_ref[idx, sel[idx]])
    assert np.array_equal(ref.total[idx, sel[idx]], ref.max_total[idx])
    assert np.array_equal(res.selected_score[idx], ref.max_total[idx])
    if per_plugin:
        pend_table = pk.pods(pending)
        for bit, raw_ref, norm_ref, w in ((FLEX, ref.raw_flex, ref.norm_flex, cfg.weight_flex), (FIT, ref.raw_fit, ref.norm_fit, cfg.weight_fit),
                                          (ALLOC, ref.raw_alloc, ref.norm_alloc, cfg.weight_alloc)):
            if not (cfg.score_mask & bit):
                continue
            r_raw = eng.evaluate(pend_table, want_planes=True, emit_mask=bit, emit_raw=1)
            assert np.array_equal(r_raw.scores[:, :N][m], raw_ref[m])
            r_norm = eng.evaluate(pend_table, want_planes=True, emit_mask=bit)
            assert np.array_equal(r_norm.scores[:, :N][m], norm_ref[m] * (w if w else 1))
    eng.close()
    return res, took, looked_up


CONFIG_CASES = [
    # (name, synth kwargs, plugin kwargs)
    ("c1-flex-allocatable", dict(cfg=1, n_nodes=700, n_pending=600), None),
    ("c2-flex-least", dict(cfg=2, n_nodes=1100, n_pending=500), None),
    ("c2-quirks-panics", dict(cfg=2, n_nodes=600, n_pending=400, pm_quirk_nodes=150, pm_panic_nodes=60, pm_nongpu_nodes=80),
     None),
    ("most-allocated-weights", dict(cfg=2, n_nodes=520, n_pending=300),
     dict(score_plugins=(("FlexGPU", 3), ("NodeResourcesFit", 2)), fit_strategy="MostAllocated", fit_resources=(("cpu", 3), ("memory", 7)))),
    ("flex-first-order", dict(cfg=2, n_nodes=520, n_pending=300, pm_panic_nodes=80),
     dict(filter_plugins=("FlexGPU", "NodeResourcesFit"))),
    ("all-three-scores", dict(cfg=2, n_nodes=513, n_pending=257),
     dict(score_plugins=(("FlexGPU", 2), ("NodeResourcesFit", 1), ("NodeResourcesAllocatable", 5)), alloc_mode="Most")),
    ("generic-fit-resources", dict(cfg=2, n_nodes=400, n_pending=300, n_scalars=4, pm_extra_scalar=400),
     dict(fit_resources=(("cpu", 2), ("memory", 1), ("ephemeral-storage", 1), (("scalar", 0), 4), (("scalar", 2), 1)))),
    ("allocatable-scalar-and-unknown", dict(cfg=1, n_nodes=300, n_pending=200),
     dict(score_plugins=(("NodeResourcesAllocatable", 2),), alloc_resources=(("cpu", 1 << 20), ("memory", 1), (("scalar", 1), 7), ("pods", 3)))),
    ("filter-only-flex", dict(cfg=2, n_nodes=300, n_pending=200), dict(filter_plugins=("FlexGPU",), score_plugins=(("FlexGPU", 1),))),
    ("ignored-scalar", dict(cfg=2, n_nodes=300, n_pending=200, n_scalars=3, pm_extra_scalar=500), dict(fit_ignored_scalar_mask=0b100)),
    ("huge-weight-keys", dict(cfg=2, n_nodes=300, n_pending=200), dict(score_plugins=(("FlexGPU", 50000), ("NodeResourcesFit", 70000)))),
    # FlexGPU scores without filtering: pods with BOTH limits stay feasible and are scored by GPUScore (flex_gpu.go:157-159)
    ("flex-score-without-flex-filter", dict(cfg=2, n_nodes=400, n_pending=300, pm_degenerate=300, pm_invalid_nodes=20),
     dict(filter_plugins=("NodeResourcesFit",), score_plugins=(("FlexGPU", 2), ("NodeResourcesFit", 1)))),
]


@pytest.mark.parametrize("name,skw,pkw", CONFIG_CASES, ids=[c[0] for c in CONFIG_CASES])
def test_evaluate_matches_oracle(gpu_ok, oracle, name, skw, pkw):
    skw = dict(skw)
    cfg_id = skw.pop("cfg")
    s = fgs.config_synth(cfg_id, **skw)
    cfg = fgs.config_plugins(cfg_id) if pkw is None else fgs.plugin_cfg(**pkw)
    ref, res = check_against_oracle(oracle, s.snapshot, s.pending, cfg, S=s.params["n_scalars"], tag=name)
    # the synthetic data must exercise what the case is about
    assert ref.feasible_count.max() > 1
    if "panic" in name:
        assert (ref.pod_err & L.FGS_PODERR_REF_PANIC).any()
    # which kernel family the unforced run took: the benchmark configurations must qualify for int32,
    # configurations outside its scope must have fallen back to int64
    expect_int32 = {"c1-flex-allocatable": (True, True), "c2-flex-least": (True, True), "c2-quirks-panics": (True, True),
                    "most-allocated-weights": (True, True), "all-three-scores": (True, True), "filter-only-flex": (True, True),
                    "generic-fit-resources": (False, False), "huge-weight-keys": (False, False)}
    if name in expect_int32:
        assert PATHS_TAKEN[name] == expect_int32[name], PATHS_TAKEN[name]


def _flex_node(B, gpus, mem_alloc, used):
    pods = [B.pod(containers=[B.ctr({("scalar", 1): u}, mem_limit=u)], node_name=True, index=i) for i, u in enumerate(used) if u]
    return B.node(alloc={"cpu": 64000, "memory": 1 << 40, "pods": 110, ("scalar", 0): gpus, ("scalar", 1): mem_alloc}, pods=pods)


def test_negative_memscore_range_error(gpu_ok, oracle):
    """The one way a normalised score leaves [0,100] on this path (runtime/framework.go:920-923): MemScore < 0
    on a feasible node. It needs memory allocatable not divisible by the gpu count, an over-committed device
    and a pod whose memory limit key is present with value 0:
      node 0: 2 gpus, 2001 memory → memEach 1000; devices used 1000 and 1001 → limit sum 2001 + 0 <= 2001 passes,
              device 0 fits (1000 >= 1000 + 0) → feasible; MemScore = 2000 - 2001 = -1
      node 1: memEach = other/2, empty → MemScore = other
    other = 100  → max 100:  100*(-1)/100 = -1 → 100 - (-1) = 101 → the cycle errors out
    other = 1000 → max 1000: 100*(-1)/1000 truncates to 0 → 100 → fine."""
    import builders as B
    cfg = fgs.plugin_cfg(filter_plugins=("FlexGPU",), score_plugins=(("FlexGPU", 1),))
    pend = B.build_podset([B.pod(containers=[B.ctr({"cpu": 100}, mem_limit=0)])])
    for other, want_err in ((100, L.FGS_PODERR_SCORE_RANGE), (1000, 0)):
        snap = B.build_snapshot([_flex_node(B, 2, 2001, [1000, 1001]), _flex_node(B, 2, other, [0, 0])])
        ref, res = check_against_oracle(oracle, snap.ptr, pend.ptr, cfg)
        assert ref.feasible_count.tolist() == [2]
        assert ref.raw_flex[0].tolist() == [-1, other]
        assert ref.pod_err.tolist() == [want_err]
        assert res.pod_err.tolist() == [want_err]


def test_whole_gpu_pod_skips_device_check(gpu_ok, oracle):
    """FlexGPU.Filter checks devices only for memory pods (flex_gpu.go:109): a whole-GPU pod fits a node whose
    devices are all partly used as long as the limit sums pass; its score is the count of idle devices."""
    import builders as B
    cfg = fgs.plugin_cfg(filter_plugins=("FlexGPU",), score_plugins=(("FlexGPU", 1),))
    snap = B.build_snapshot([_flex_node(B, 2, 2000, [500, 700]), _flex_node(B, 2, 2000, [0, 700]), _flex_node(B, 2, 2000, [0, 0])])
    pend = B.build_podset([B.pod(containers=[B.ctr({"cpu": 100}, gpu_limit=1)]), B.pod(containers=[B.ctr({"cpu": 100}, gpu_limit=3)])])
    ref, res = check_against_oracle(oracle, snap.ptr, pend.ptr, cfg)
    assert ref.feasible_count.tolist() == [3, 0]  # limit 3 > 2 allocatable everywhere
    assert ref.raw_flex[0].tolist() == [0, 1, 2]
    assert ref.norm_flex[0].tolist() == [100, 50, 0]  # reverse normalisation: bin-packing prefers the fuller node


def test_edge_shapes(gpu_ok, oracle):
    """ragged sizes around the tile boundaries, a single node, a single pod"""
    for n_nodes, n_pending in ((1, 1), (31, 65), (512, 64), (513, 63), (1025, 129)):
        s = fgs.config_synth(2, n_nodes=n_nodes, n_pending=n_pending)
        check_against_oracle(oracle, s.snapshot, s.pending, fgs.config_plugins(2), per_plugin=False)


def test_empty_pod_batch(gpu_ok):
    s = fgs.config_synth(2, n_nodes=100, n_pending=0)
    eng, pk, res = run_engine((s.snapshot, s.pending), fgs.config_plugins(2), want_planes=False)
    assert res.feasible_count.shape == (0,)
    eng.close()


def test_plugin_arg_validation(gpu_ok):
    # pkg/noderesources/allocatable_test.go:220-235 and NewAllocatable's mode check (allocatable.go:93-98)
    eng = fgs.Engine(device=0)
    for case in [c for c in G.load("allocatable") if "want_err" in c]:
        cfg = fgs.plugin_cfg(score_plugins=(("NodeResourcesAllocatable", 1),), alloc_resources=[tuple(r) for r in case["resources"]])
        with pytest.raises(fgs.FgsError) as ei:
            eng.set_plugins(cfg)
        assert case["want_err"] in str(ei.value)
    with pytest.raises(fgs.FgsError):
        eng.set_plugins(fgs.plugin_cfg(score_plugins=(("NodeResourcesAllocatable", 1),), alloc_mode="Sideways"))
    eng.close()


@pytest.mark.parametrize("case", [c for c in G.load("allocatable") if "want_err" not in c], ids=lambda c: c["name"][:60])
def test_golden_allocatable(gpu_ok, case):
    # pkg/noderesources/allocatable_test.go:112-219 through the CUDA path
    snap, pend, resources, mode = G.allocatable_case(case)
    cfg = fgs.plugin_cfg(filter_plugins=(), score_plugins=(("NodeResourcesAllocatable", 1),), alloc_mode=mode, alloc_resources=resources)
    eng, pk, res = run_engine((snap.ptr, pend.ptr), cfg)
    n = len(case["expected"])
    assert res.scores[0, :n].tolist() == case["expected"]
    eng.close()


def test_golden_gang(gpu_ok, oracle):
    # pkg/coscheduling/core/core_test.go:42-234 through the CUDA path. The evaluated pod is part of the pending
    # set here (in production the pod lister knows it), which does not change any expected outcome.
    g = G.load("gang_prefilter")
    for case in g["cases"]:
        snap, pend, pgs, unlisted, me = G.gang_prefilter_case(g, case)
        eng = fgs.Engine(device=0)
        eng.upload_nodes(fgs.Packer(2).nodes(snap.ptr))
        pre, permit, assigned = eng.gang_check(snap.ptr, pend.ptr, pgs.ptr)
        assert (pre[me] == 0) == case["expected"], case["name"]
        assert np.array_equal(pre, oracle.gang_prefilter(snap.ptr, pend.ptr, pgs.ptr))
        eng.close()
    g = G.load("gang_permit")
    for case in g["cases"]:
        snap, pend, pgs = G.gang_permit_case(g, case)
        eng = fgs.Engine(device=0)
        eng.upload_nodes(fgs.Packer(2).nodes(snap.ptr))
        pre, permit, assigned = eng.gang_check(snap.ptr, pend.ptr, pgs.ptr)
        assert int(permit[0]) == case["expected"], case["name"]
        eng.close()


def test_golden_capacity(gpu_ok):
    # pkg/capacityscheduling/capacity_scheduling_test.go:52-164 through the CUDA path
    fw = {0: 0, 1: 2, 2: 2}
    for case in G.load("capacity_prefilter"):
        pend, qs = G.capacity_case(case)
        eng = fgs.Engine(device=0, n_scalars=3)
        code = eng.quota_check(fgs.Packer(3).pods(pend.ptr), qs.ptr)
        assert [fw[int(c)] for c in code] == case["expected"], case["name"]
        eng.close()


@pytest.mark.parametrize("seed", [0, 1])
def test_gang_matches_oracle(gpu_ok, oracle, seed):
    s = fgs.config_synth(3, n_nodes=1500, n_pending=4000, n_groups=150, n_namespaces=7, pm_invalid_nodes=30, pm_neg_left_nodes=60,
                         seed_nodes=0xF6500003 + seed, seed_pods=0xF6510003 + seed)
    eng = fgs.Engine(device=0)
    eng.upload_nodes(fgs.Packer(2).nodes(s.snapshot))
    pre, permit, assigned = eng.gang_check(s.snapshot, s.pending, s.podgroups)
    ref_pre = oracle.gang_prefilter(s.snapshot, s.pending, s.podgroups)
    ref_permit, ref_assigned = oracle.gang_permit(s.snapshot, s.pending, s.podgroups)
    assert np.array_equal(pre, ref_pre)
    assert np.array_equal(permit, ref_permit)
    assert np.array_equal(assigned, ref_assigned)
    assert set(np.unique(ref_pre)) >= {0, 2, 3}  # the generator reaches every outcome
    eng.close()


@pytest.mark.parametrize("n_nominated", [0, 40])
def test_quota_matches_oracle(gpu_ok, oracle, n_nominated):
    s = fgs.config_synth(4, n_nodes=800, n_pending=5000, n_namespaces=40, n_nominated=n_nominated, n_scalars=3, pm_extra_scalar=300)
    eng = fgs.Engine(device=0, n_scalars=3)
    pk_p, pk_n = fgs.Packer(3), fgs.Packer(3)
    P = s.n_pending
    puid = np.arange(P, dtype=np.int64)
    nuid = np.arange(n_nominated, dtype=np.int64) * 7  # some nominated pods ARE pending pods (same uid)
    nom_t = pk_n.pods(s.nominated) if n_nominated else None
    code = eng.quota_check(pk_p.pods(s.pending), s.quotas, nom_t, puid if n_nominated else None, nuid if n_nominated else None)
    ref = oracle.quota_prefilter(s.pending, s.quotas, s.nominated if n_nominated else None, puid if n_nominated else None,
                                 nuid if n_nominated else None)
    assert np.array_equal(code, ref)
    assert set(np.unique(ref)) >= {0, 1}  # the "over Σmin" outcome (2) is pinned by the golden vectors
    eng.close()


def test_dirty_row_update(gpu_ok, oracle):
    """fgs_update_nodes on a few rows == a full upload of the changed snapshot"""
    a = fgs.config_synth(2, n_nodes=600, n_pending=200)
    b = fgs.config_synth(2, n_nodes=600, n_pending=200, seed_nodes=0xABCDEF)  # same pods, different nodes
    cfg = fgs.config_plugins(2)
    eng = fgs.Engine(device=0)
    eng.set_plugins(cfg)
    pk, pk2, pkp = fgs.Packer(2), fgs.Packer(2), fgs.Packer(2)
    eng.upload_nodes(pk.nodes(a.snapshot))
    dirty = np.array([3, 77, 511, 512, 599], np.int64)
    # rows taken from b replace the same indices of a: pack b's single-node shards
    import ctypes as C
    full_b = pk2.nodes(b.snapshot)
    N = 600
    for i in dirty:
        row = fgs.Packer(2)
        eng.update_nodes(row.nodes(b.snapshot, int(i), int(i) + 1), [int(i)])
    res = eng.evaluate(pkp.pods(a.pending), want_planes=True)
    # reference: oracle on a with those rows replaced — build by evaluating both and splicing columns
    ra = oracle.evaluate(a.snapshot, a.pending, cfg, workers=4)
    rb = oracle.evaluate(b.snapshot, a.pending, cfg, workers=4)
    feas = res.feasible(N)
    exp = ra.status == 0
    exp[:, dirty] = (rb.status == 0)[:, dirty]
    assert np.array_equal(feas, exp)
    eng.close()


def test_reason_planes_match_the_oracle(gpu_ok, oracle):
    """fgs_reason_plane: the message inside each Filter status (flex_gpu.go:60-114) and Fit's InsufficientResource
    list (fit.go:256-329), against the oracle's walk."""
    s = fgs.config_synth(2, n_nodes=900, n_pending=96, pm_panic_nodes=15, pm_nongpu_nodes=60, pm_extra_scalar=200, n_scalars=3)
    cfg = fgs.config_plugins(2)
    eng = fgs.Engine(device=0, n_scalars=3)
    eng.set_plugins(cfg)
    eng.upload_nodes(fgs.Packer(3).nodes(s.snapshot))
    eng.evaluate(fgs.Packer(3).pods(s.pending))
    flex = eng.reason_plane(L.FGS_PL_FLEXGPU)
    fit = eng.reason_plane(L.FGS_PL_FIT)
    seen_flex, seen_fit = set(), set()
    for p in range(s.n_pending):
        rf, rt = oracle.filter_reasons(s.snapshot, s.pending, cfg, p)
        assert np.array_equal(flex[p], rf), p
        assert np.array_equal(fit[p], rt), p
        seen_flex |= set(int(x) for x in rf)
        seen_fit |= set(int(x) for x in rt)
    assert {0, 1, 2, 3, 4, 5, 0xF1} <= seen_flex, seen_flex
    # cpu, memory and all three scalar slots occur, alone and combined
    assert all(any(x & b for x in seen_fit) for b in (2, 4, 16, 32, 64)) and any(bin(x).count("1") >= 3 for x in seen_fit)
    eng.close()
    # "Too many pods" (fit.go:259-268), alone and together with another dimension
    full = B.node(alloc={"cpu": 4000, "memory": 8 << 30, "pods": 1}, pods=[B.pod([B.ctr({"cpu": 100})], node_name=True)])
    free = B.node(alloc={"cpu": 4000, "memory": 8 << 30, "pods": 110})
    snap = B.build_snapshot([full, free])
    pend = B.build_podset([B.pod([B.ctr({"cpu": 100})]), B.pod([B.ctr({"cpu": 8000})]), B.pod([B.ctr({})])])
    eng = fgs.Engine(device=0)
    eng.set_plugins(cfg)
    eng.upload_nodes(fgs.Packer(2).nodes(snap.ptr))
    eng.evaluate(fgs.Packer(2).pods(pend.ptr))
    fit = eng.reason_plane(L.FGS_PL_FIT)
    assert fit.tolist() == [[1, 0], [3, 2], [1, 0]]
    for p in range(3):
        assert np.array_equal(fit[p], oracle.filter_reasons(snap.ptr, pend.ptr, cfg, p)[1])
    eng.close()


def test_submit_collect_pipeline_equals_evaluate(gpu_ok):
    """fgs_submit / fgs_collect (two batches in flight, double-buffered staging) returns what fgs_evaluate returns,
    batch after batch, including batches of different sizes."""
    s = fgs.config_synth(2, n_nodes=3000, n_pending=5000, pm_extra_scalar=0)
    cfg = fgs.config_plugins(2)
    eng = fgs.Engine(device=0)
    eng.set_plugins(cfg)
    eng.upload_nodes(fgs.Packer(2).nodes(s.snapshot))
    bounds = [(0, 700), (700, 2048), (2048, 2049), (2049, 4000), (4000, 5000), (0, 5000), (100, 164)]
    packers = [fgs.Packer(2) for _ in bounds]
    tables = [pk.pods(s.pending, a, b) for pk, (a, b) in zip(packers, bounds)]
    want = [eng.evaluate(t) for t in tables]
    got = []
    eng.submit(tables[0])
    for t in tables[1:]:
        eng.submit(t)
        got.append(eng.collect())
    got.append(eng.collect())
    for w, g in zip(want, got):
        assert np.array_equal(w.feasible_count, g.feasible_count)
        assert np.array_equal(w.selected_node, g.selected_node)
        assert np.array_equal(w.selected_score, g.selected_score)
        assert np.array_equal(w.pod_err, g.pod_err)
    # a third submit without a collect is refused, and so is the synchronous path while batches are in flight
    eng.submit(tables[0])
    eng.submit(tables[1])
    with pytest.raises(fgs.FgsError):
        eng.submit(tables[2])
    with pytest.raises(fgs.FgsError):
        eng.evaluate(tables[2])
    a, b = eng.collect(), eng.collect()
    assert np.array_equal(a.selected_node, want[0].selected_node) and np.array_equal(b.selected_node, want[1].selected_node)
    again = eng.evaluate(tables[3])  # the synchronous path works again once everything is collected
    assert np.array_equal(again.selected_node, want[3].selected_node)
    eng.close()


def _expected_fit_error_histogram(oracle, snapshot, pending, cfg, ref, n_nodes, node_valid):
    """FitError's reason counts from the oracle's walk: per unschedulable pod, every node's status is the status of the
    first filter plugin (profile order) that rejected it"""
    order = [int(x) for x in cfg.filter_order if int(x)]
    exp = np.zeros((len(ref.feasible_count), L.FGS_DIAG_SLOTS), np.int32)
    for p in np.nonzero(ref.feasible_count == 0)[0]:
        rf, rt = oracle.filter_reasons(snapshot, pending, cfg, int(p))
        for n in range(n_nodes):
            if not node_valid[n]:
                exp[p, L.FGS_DIAG_NOT_A_NODE] += 1
                continue
            for pl in order:
                if pl == L.FGS_PL_FLEXGPU and rf[n]:
                    exp[p, L.FGS_DIAG_FLEX_PANIC if rf[n] == 0xF1 else L.FGS_DIAG_FLEX0 + int(rf[n]) - 1] += 1
                    break
                if pl == L.FGS_PL_FIT and rt[n]:
                    for b in range(16):
                        if (int(rt[n]) >> b) & 1:
                            exp[p, L.FGS_DIAG_FIT0 + b] += 1
                    break
    return exp


def test_fit_error_histogram_matches_the_oracle(gpu_ok, oracle):
    """fgs_fit_error_histogram: the counts FitError.Error() prints (framework/types.go:221-250), per unschedulable pod"""
    for order in (("NodeResourcesFit", "FlexGPU"), ("FlexGPU", "NodeResourcesFit")):
        s = fgs.config_synth(2, n_nodes=300, n_pending=400, pm_panic_nodes=15, pm_nongpu_nodes=200, pm_extra_scalar=200, pm_invalid_nodes=20,
                             n_scalars=3, max_existing_pods=60)
        cfg = fgs.plugin_cfg(filter_plugins=order, score_plugins=(("FlexGPU", 1), ("NodeResourcesFit", 1)))
        eng = fgs.Engine(device=0, n_scalars=3)
        eng.set_plugins(cfg)
        eng.upload_nodes(fgs.Packer(3).nodes(s.snapshot))
        res = eng.evaluate(fgs.Packer(3).pods(s.pending))
        got = eng.fit_error_histogram()
        ref = oracle.evaluate(s.snapshot, s.pending, cfg, workers=4)
        assert np.array_equal(res.feasible_count, ref.feasible_count)
        valid = np.ctypeslib.as_array(s.snapshot.contents.node_valid, shape=(s.n_nodes,))
        exp = _expected_fit_error_histogram(oracle, s.snapshot, s.pending, cfg, ref, s.n_nodes, valid)
        assert int((ref.feasible_count == 0).sum()) >= 10
        assert np.array_equal(got, exp), np.nonzero((got != exp).any(axis=1))[0][:5]
        used = set(np.nonzero(exp.sum(axis=0))[0].tolist())
        assert {L.FGS_DIAG_FLEX0, L.FGS_DIAG_NOT_A_NODE} <= used and len(used) >= 5, used
        eng.close()
    # by hand: 3 nodes, nothing fits. Fit first: node 0 lacks cpu AND memory (both are listed), node 1 cpu, node 2 has
    # no flex resources but Fit already said no for its pod count
    GiB = 1 << 30
    nodes = [B.node({"cpu": 1000, "memory": GiB, "pods": 110, ("scalar", 0): 8, ("scalar", 1): 64 * GiB}),
             B.node({"cpu": 1000, "memory": 64 * GiB, "pods": 110, ("scalar", 0): 8, ("scalar", 1): 64 * GiB}),
             B.node({"cpu": 64000, "memory": 64 * GiB, "pods": 1}, [B.pod([B.ctr({"cpu": 1})], node_name=True)])]
    pend = [B.pod([B.ctr({"cpu": 2000, "memory": 2 * GiB}, mem_limit=GiB)]), B.pod([B.ctr({"cpu": 10})])]
    snap, ps = B.build_snapshot(nodes), B.build_podset(pend)
    for order, want in ((("NodeResourcesFit", "FlexGPU"), {L.FGS_DIAG_FIT0 + 1: 2, L.FGS_DIAG_FIT0 + 2: 1, L.FGS_DIAG_FIT0 + 0: 1}),
                        (("FlexGPU", "NodeResourcesFit"), {L.FGS_DIAG_FIT0 + 1: 2, L.FGS_DIAG_FIT0 + 2: 1, L.FGS_DIAG_FLEX0 + 1: 1})):
        cfg = fgs.plugin_cfg(filter_plugins=order, score_plugins=(("NodeResourcesFit", 1),))
        eng = fgs.Engine(device=0)
        eng.set_plugins(cfg)
        eng.upload_nodes(fgs.Packer(2).nodes(snap.ptr))
        res = eng.evaluate(fgs.Packer(2).pods(ps.ptr))
        assert res.feasible_count.tolist() == [0, 2]
        got = eng.fit_error_histogram()
        assert {k: int(v) for k, v in enumerate(got[0]) if v} == want, (order, got[0])
        assert not got[1].any()  # a pod some node took has no FitError
        eng.close()


def test_fused_pass_equals_the_two_passes(gpu_ok, oracle, monkeypatch):
    """Filter + Score in one pass (flexmax prepass + fused kernel) gives bit for bit what reduce + emit give — bitmap,
    counts, error bits, every score, the selected node — over plugin sets, strategies, filter orders and quirk nodes;
    and both equal the oracle."""
    s = fgs.config_synth(2, n_nodes=1700, n_pending=900, pm_panic_nodes=20, pm_nongpu_nodes=80, pm_invalid_nodes=15, pm_quirk_nodes=40,
                         max_existing_pods=40)
    variants = [
        dict(),
        dict(filter_plugins=("FlexGPU", "NodeResourcesFit")),
        dict(filter_plugins=("FlexGPU",), score_plugins=(("FlexGPU", 3),)),
        dict(filter_plugins=("NodeResourcesFit",), score_plugins=(("NodeResourcesFit", 2),)),
        dict(fit_strategy="MostAllocated", score_plugins=(("FlexGPU", 2), ("NodeResourcesFit", 5))),
        dict(fit_resources=(("cpu", 3), ("memory", 1))),
        dict(filter_plugins=("NodeResourcesFit",), score_plugins=(("FlexGPU", 1), ("NodeResourcesFit", 1))),
    ]
    for kw in variants:
        cfg = fgs.plugin_cfg(**kw)
        out = []
        for no_fused in ("0", "1"):
            monkeypatch.setenv("FGS_FUSED", "1" if no_fused == "0" else "0")
            eng = fgs.Engine(device=0)
            eng.set_plugins(cfg)
            eng.upload_nodes(fgs.Packer(2).nodes(s.snapshot))
            r = eng.evaluate(fgs.Packer(2).pods(s.pending), want_planes=True)
            t = eng.timings()
            assert t["fused"] == (no_fused == "0"), (kw, t)
            assert t["emit_int32"]
            out.append(r)
            eng.close()
        a, b = out
        assert np.array_equal(a.feasible_bitmap, b.feasible_bitmap), kw
        assert np.array_equal(a.feasible_count, b.feasible_count), kw
        assert np.array_equal(a.pod_err, b.pod_err), kw
        assert np.array_equal(a.selected_node, b.selected_node) and np.array_equal(a.selected_score, b.selected_score), kw
        feas = a.feasible(s.n_nodes)
        assert np.array_equal(a.scores[:, :s.n_nodes][feas], b.scores[:, :s.n_nodes][feas]), kw
        ref = oracle.evaluate(s.snapshot, s.pending, cfg, workers=4)
        assert np.array_equal(feas, ref.status == 0), kw
        assert np.array_equal(a.feasible_count, ref.feasible_count) and np.array_equal(a.pod_err, ref.pod_err), kw
        ok = (ref.pod_err == 0) & (ref.feasible_count >= 1)
        m = feas & ok[:, None]
        assert np.array_equal(a.scores[:, :s.n_nodes][m], ref.total[m]), kw
    monkeypatch.delenv("FGS_FUSED")


@pytest.mark.parametrize("cfg_id,S,extra", [(2, 2, {}), (1, 2, {}), (2, 4, dict(pm_extra_scalar=400))])
def test_pinned_packer_tables_go_by_dma_and_give_the_same_results(gpu_ok, oracle, cfg_id, S, extra):
    """fgs_packer_pin + fgs_pod_stats: the engine copies the columns straight from the packer's table (no staging pass);
    results equal those of the staged path bit for bit, in the blocking call and in the submit / collect pipeline"""
    s = fgs.config_synth(cfg_id, n_nodes=900, n_pending=3000, n_scalars=S, **extra)
    cfg = fgs.config_plugins(cfg_id)
    eng = fgs.Engine(device=0, n_scalars=S)
    eng.set_plugins(cfg)
    eng.upload_nodes(fgs.Packer(S).nodes(s.snapshot))
    bounds = [(0, 3000), (5, 1030), (1030, 1031), (2000, 2960)]
    plain = [fgs.Packer(S) for _ in bounds]
    pinned = [fgs.Packer(S, pin=True) for _ in bounds]
    want = [eng.evaluate(pk.pods(s.pending, a, b), want_planes=True) for pk, (a, b) in zip(plain, bounds)]
    tabs = [pk.pods(s.pending, a, b) for pk, (a, b) in zip(pinned, bounds)]
    for t in tabs:
        st = t.contents.stats.contents
        assert st.flags & 1 and st.max_val[2] > 0
    N = s.n_nodes
    for w, t in zip(want, tabs):
        g = eng.evaluate(t, want_planes=True)
        assert np.array_equal(w.feasible(N), g.feasible(N))
        f = w.feasible(N)
        assert np.array_equal(w.scores[:, :N][f], g.scores[:, :N][f])
        for k in ("feasible_count", "selected_node", "selected_score", "pod_err"):
            assert np.array_equal(getattr(w, k), getattr(g, k)), k
    got = []
    eng.submit(tabs[0])
    for t in tabs[1:]:
        eng.submit(t)
        got.append(eng.collect())
    got.append(eng.collect())
    for w, g in zip(want, got):
        for k in ("feasible_count", "selected_node", "selected_score", "pod_err"):
            assert np.array_equal(getattr(w, k), getattr(g, k)), k
    # the oracle on the first batch, so that the DMA path stands on its own
    ref = oracle.evaluate(s.snapshot, s.pending, cfg, 0, 3000, workers=4)
    assert np.array_equal(got[0].feasible_count, ref.feasible_count) and np.array_equal(got[0].pod_err, ref.pod_err)
    ok = (ref.pod_err == 0) & (ref.feasible_count > 0)
    assert np.array_equal(got[0].selected_score[ok], ref.max_total[ok])
    eng.close()


def test_resident_batches_and_resubmit_on_one_engine(gpu_ok):
    """fgs_set_resident_batches + fgs_resubmit: a job of five batches stays resident in five pipeline slots; walking the
    ring again (two batches in flight) returns each batch's results in order; a node update makes the resident batches
    stale"""
    s = fgs.config_synth(2, n_nodes=1500, n_pending=2600)
    cfg = fgs.config_plugins(2)
    eng = fgs.Engine(device=0)
    eng.set_plugins(cfg)
    pkn = fgs.Packer(2)
    eng.upload_nodes(pkn.nodes(s.snapshot))
    bounds = [(0, 640), (640, 1000), (1000, 1001), (1001, 2000), (2000, 2600)]
    packers = [fgs.Packer(2) for _ in bounds]
    tables = [pk.pods(s.pending, a, b) for pk, (a, b) in zip(packers, bounds)]
    want = [eng.evaluate(t) for t in tables]
    eng.set_resident_batches(len(tables))

    def walk(sub):
        out, inflight = [], 0
        for i in range(len(tables)):
            sub(i)
            inflight += 1
            if inflight == 2:
                out.append(eng.collect())
                inflight -= 1
        while inflight:
            out.append(eng.collect())
            inflight -= 1
        return out

    first = walk(lambda i: eng.submit(tables[i]))
    for _ in range(2):
        again = walk(lambda i: eng.resubmit(bounds[i][1] - bounds[i][0]))
        for w, a, b in zip(want, first, again):
            for k in ("feasible_count", "selected_node", "selected_score", "pod_err"):
                assert np.array_equal(getattr(w, k), getattr(a, k)) and np.array_equal(getattr(w, k), getattr(b, k)), k
    eng.update_nodes(pkn.nodes(s.snapshot, 3, 4), [3])
    with pytest.raises(fgs.FgsError):
        eng.resubmit(640)
    with pytest.raises(fgs.FgsError):
        eng.set_resident_batches(33)
    eng.set_resident_batches(0)
    r = eng.evaluate(tables[1])
    assert np.array_equal(r.feasible_count, want[1].feasible_count)
    eng.close()


def test_gang_postfilter_decision_around_the_ten_percent_gap(gpu_ok):
    """Coscheduling.PostFilter (pkg/coscheduling/coscheduling.go:140-176): assigned >= MinMember keeps the group, a gap of
    at most 10 % — float32(MinMember - assigned) / float32(MinMember) <= 0.1 in single precision, as Go computes it —
    keeps it too, anything else rejects the whole group. Hand cases on both sides of the edge, plus a synthetic
    cluster against the same formula in numpy float32."""
    cases = [(10, 9), (10, 8), (11, 10), (11, 9), (20, 18), (20, 17), (9, 8), (30, 27), (30, 26), (7, 7), (5, 6), (1000, 900), (1000, 899),
             (41, 37), (3, 0), (1, 0)]
    groups = [dict(min_member=mm, ns=0) for mm, _ in cases] + [dict(min_member=4, ns=0, exists=False)]
    on_node = []
    for g, (mm, a) in enumerate(cases):
        on_node += [B.pod([B.ctr({"cpu": 10})], pg=g, ns=0, node_name=True) for _ in range(a)]
    on_node += [B.pod([B.ctr({"cpu": 10})], pg=0, ns=0, node_name=False)]  # not assigned: Spec.NodeName is empty (core.go:311)
    on_node += [B.pod([B.ctr({"cpu": 10})], pg=1, ns=3, node_name=True)]   # another namespace: not a member
    snap = B.build_snapshot([B.node({"cpu": 10 ** 7, "memory": 1 << 40, "pods": 10 ** 5}, on_node)])
    pend = B.build_podset([B.pod([B.ctr({"cpu": 10})], pg=g, ns=0) for g in range(len(groups))] + [B.pod([B.ctr({"cpu": 10})], pg=-1, ns=0)])
    pgs = B.build_podgroups(groups)
    eng = fgs.Engine(device=0)
    eng.upload_nodes(fgs.Packer(2).nodes(snap.ptr))
    code = eng.gang_postfilter(snap.ptr, pend.ptr, pgs.ptr)
    want = []
    for mm, a in cases:
        if a >= mm:
            want.append(2)
        else:
            want.append(3 if np.float32(mm - a) / np.float32(mm) <= np.float32(0.1) else 4)
    want += [1, 1]  # PodGroup object not found; pod without the label
    assert list(code) == want, (list(code), want)
    assert want[:9] == [3, 4, 3, 4, 3, 4, 4, 3, 4]  # the edge itself, written out: 1/10, 2/20 and 3/30 pass, 1/9 and 3/20 do not
    eng.close()
    s = fgs.config_synth(3, n_nodes=800, n_pending=3000, n_groups=120, n_namespaces=5)
    eng = fgs.Engine(device=0)
    eng.upload_nodes(fgs.Packer(2).nodes(s.snapshot))
    code = eng.gang_postfilter(s.snapshot, s.pending, s.podgroups)
    _, _, assigned = eng.gang_check(s.snapshot, s.pending, s.podgroups)
    G_ = int(s.podgroups.contents.n_groups)
    pg = np.ctypeslib.as_array(s.pending.contents.pg, shape=(s.n_pending,))
    mm = np.ctypeslib.as_array(s.podgroups.contents.min_member, shape=(G_,))
    ex = np.ctypeslib.as_array(s.podgroups.contents.exists, shape=(G_,))
    exp = np.ones(s.n_pending, np.uint8)
    has = pg >= 0
    g = pg[has]
    a = assigned[g]
    gap = ((mm[g] - a).astype(np.float32) / mm[g].astype(np.float32)) <= np.float32(0.1)
    exp[has] = np.where(ex[g] == 0, 1, np.where(a >= mm[g], 2, np.where(gap, 3, 4)))
    assert np.array_equal(code, exp)
    assert 4 in set(np.unique(code)) and len(set(np.unique(code))) >= 2
    eng.close()


def test_sync_nodes_repacks_only_the_rows_whose_generation_moved(gpu_ok, oracle):
    """fgs_sync_nodes: the device tables follow the snapshot the way Cache.UpdateSnapshot follows the cache
    (vendor/k8s.io/kubernetes/pkg/scheduler/internal/cache/cache.go:197-276) — by NodeInfo.Generation. After every sync
    the evaluation equals the oracle's on the NEW snapshot and that of an engine that uploaded everything."""
    import random
    from oracle import cross_fuzz as F
    rng = random.Random(99)
    S = 2
    nodes = [F.rand_node(rng, S) for _ in range(300)]
    for nd in nodes:  # at most 8 devices at first, so that a later 11-device row overflows the uploaded slots
        if nd["alloc"].get(("scalar", 0), 0) > 8:
            nd["alloc"][("scalar", 0)] = 8
    pods = [F.rand_pod(rng, S) for _ in range(200)]
    pend = B.build_podset(pods, S)
    cfg = fgs.config_plugins(2)
    eng = fgs.Engine(device=0)
    eng.set_plugins(cfg)
    pk = fgs.Packer(S)
    table = fgs.Packer(S).pods(pend.ptr)
    gen = np.arange(1, len(nodes) + 1, dtype=np.int64)
    next_gen = [len(nodes) + 1]

    def sync_and_check(expect_repacked):
        snap = B.build_snapshot(nodes, S)
        snap.struct.generation = gen.ctypes.data_as(L.p_i64)
        n = eng.sync_nodes(pk, snap.ptr)
        assert n == expect_repacked, (n, expect_repacked)
        got = eng.evaluate(table, want_planes=True)
        ref = oracle.evaluate(snap.ptr, pend.ptr, cfg, workers=4)
        N = len(nodes)
        assert np.array_equal(got.feasible(N), ref.status == 0)
        assert np.array_equal(got.feasible_count, ref.feasible_count) and np.array_equal(got.pod_err, ref.pod_err)
        ok = (ref.pod_err == 0) & (ref.feasible_count >= 1)
        m = (ref.status == 0) & ok[:, None]
        assert np.array_equal(got.scores[:, :N][m], ref.total[m])
        fresh = fgs.Engine(device=0)
        fresh.set_plugins(cfg)
        fresh.upload_nodes(fgs.Packer(S).nodes(snap.ptr))
        want = fresh.evaluate(table)
        assert np.array_equal(want.selected_score[ok], got.selected_score[ok]) and np.array_equal(want.feasible_count, got.feasible_count)
        fresh.close()

    def touch(i, node):
        nodes[i] = node
        gen[i] = next_gen[0]
        next_gen[0] += 1

    sync_and_check(300)          # the first call uploads everything
    sync_and_check(0)            # nothing moved
    for i in (3, 77, 299):       # three NodeInfos changed (pods added / removed, allocatable edited)
        nd = F.rand_node(rng, S)
        if nd["alloc"].get(("scalar", 0), 0) > 8:
            nd["alloc"][("scalar", 0)] = 8
        touch(i, nd)
    sync_and_check(3)
    gen[5] = next_gen[0]         # a generation bump without a change is repacked too (the cache cannot tell)
    next_gen[0] += 1
    sync_and_check(1)
    big = F.rand_node(rng, S)
    big["alloc"][("scalar", 0)] = 11   # more devices than the uploaded table has slots: everything goes up again
    big["alloc"][("scalar", 1)] = 11 * 16384
    touch(8, big)
    sync_and_check(300)
    for i in range(0, 300, 2):   # more than half stale: a full upload is cheaper than 150 patches
        gen[i] = next_gen[0]
        next_gen[0] += 1
    touch(1, F.rand_node(rng, S))
    sync_and_check(300)
    snap = B.build_snapshot(nodes, S)  # no generation column: refused
    with pytest.raises(fgs.FgsError):
        eng.sync_nodes(pk, snap.ptr)
    eng.close()


def test_calls_that_would_corrupt_batches_in_flight_are_refused(gpu_ok):
    """While submitted batches are in flight the node table, the plugin set and the nominated pods must not change (the
    emit pass of a batch may still be queued behind the call), and a second thread entering the engine gets FGS_E_STATE
    instead of sharing its staging slots."""
    import threading
    s = fgs.config_synth(2, n_nodes=2000, n_pending=20000)
    cfg = fgs.config_plugins(2)
    eng = fgs.Engine(device=0)
    eng.set_plugins(cfg)
    pkn = fgs.Packer(2)
    nodes = pkn.nodes(s.snapshot)
    eng.upload_nodes(nodes)
    t = fgs.Packer(2).pods(s.pending)
    want = eng.evaluate(t)
    eng.submit(t)
    for call in (lambda: eng.upload_nodes(nodes), lambda: eng.update_nodes(pkn.nodes(s.snapshot, 3, 4), [3]), lambda: eng.set_plugins(cfg),
                 lambda: eng.set_nominated(t, np.zeros(20000, np.int32)), lambda: eng.schedule_sequential(t), lambda: eng.evaluate(t),
                 lambda: eng.set_resident_batches(8)):
        with pytest.raises(fgs.FgsError) as ei:
            call()
        assert ei.value.code == L.FGS_E_STATE
    got = eng.collect()
    assert np.array_equal(got.selected_node, want.selected_node)
    # two threads: one evaluates in a loop, the other keeps calling in; every call either runs alone or is refused
    stop = threading.Event()
    refused, ran, wrong = [0], [0], [0]

    def intruder():
        while not stop.is_set():
            try:
                r = eng.evaluate(t)
                ran[0] += 1
                if not np.array_equal(r.selected_node, want.selected_node):
                    wrong[0] += 1
            except fgs.FgsError as ex:
                if ex.code == L.FGS_E_STATE:
                    refused[0] += 1
                else:
                    raise

    th = threading.Thread(target=intruder)
    th.start()
    mine = 0
    for _ in range(60):
        try:
            r = eng.evaluate(t)
            mine += 1
            assert np.array_equal(r.selected_node, want.selected_node)
        except fgs.FgsError as ex:
            assert ex.code == L.FGS_E_STATE
    stop.set()
    th.join()
    assert wrong[0] == 0 and mine + ran[0] > 0
    eng.close()

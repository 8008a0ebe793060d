"""fgs_update_nodes / fgs_sync_nodes for a few rows: the row-patch kernel (patch_nodes) stores, derives, narrows and
checks the rows on the device and the call returns without waiting. What the cache's incremental snapshot update
(vendor/k8s.io/kubernetes/pkg/scheduler/internal/cache/cache.go:197-276) feeds the next cycle must be what a full upload
would have fed it — against the oracle on the edited snapshot, and against an engine that re-derives everything
(FGS_PATCH=0)."""
import ctypes as C
import os

import numpy as np
import pytest

import flex_gpu_scheduler_b200 as fgs
from flex_gpu_scheduler_b200 import _lib as L

pytestmark = pytest.mark.gpu


def _engine(cfg, patch=True, S=2):
    if not patch:
        os.environ["FGS_PATCH"] = "0"
    try:
        eng = fgs.Engine(device=0, n_scalars=S)
    finally:
        os.environ.pop("FGS_PATCH", None)
    eng.set_plugins(cfg)
    return eng


def _same(a, b, N):
    for k in ("feasible_count", "selected_node", "selected_score", "pod_err"):
        assert np.array_equal(getattr(a, k), getattr(b, k)), k
    assert np.array_equal(a.feasible(N), b.feasible(N))
    ok = (a.pod_err == 0) & (a.feasible_count >= 1)
    m = a.feasible(N) & ok[:, None]
    assert np.array_equal(a.scores[:, :N][m], b.scores[:, :N][m])


@pytest.mark.parametrize("cfg_id", [2, 1])
def test_row_patches_equal_a_full_rederive(gpu_ok, oracle, cfg_id):
    """Rows of another cluster replace rows of the uploaded one, a few at a time (as kernel arguments up to four, as a
    compact buffer beyond), back and forth; quirk, panic and non-GPU nodes make rows enter and leave the special list and
    the node analysis. After every patch a small batch (one launch) and a large one equal the engine that re-derives the
    table, and the feasible sets equal the oracle's, column by column from the cluster each row came from."""
    N, P = 700, 300
    kw = dict(n_nodes=N, n_pending=P, pm_quirk_nodes=120, pm_panic_nodes=60, pm_nongpu_nodes=80, pm_invalid_nodes=30)
    a = fgs.config_synth(cfg_id, **kw)
    b = fgs.config_synth(cfg_id, seed_nodes=0xABCDEF, **kw)
    cfg = fgs.config_plugins(cfg_id)
    fast, slow = _engine(cfg), _engine(cfg, patch=False)
    pk = fgs.Packer(2)
    for e in (fast, slow):
        e.upload_nodes(pk.nodes(a.snapshot))
    ra = oracle.evaluate(a.snapshot, a.pending, cfg, workers=4)
    rb = oracle.evaluate(b.snapshot, a.pending, cfg, workers=4)
    src = np.zeros(N, np.int8)  # which cluster every row holds now
    rng = np.random.default_rng(77)
    big, small_t = fgs.Packer(2).pods(a.pending), None
    snaps = (a.snapshot, b.snapshot)
    for step in range(14):
        k = int(rng.choice([1, 1, 2, 4, 5, 9, 40]))
        rows = np.sort(rng.choice(N, size=k, replace=False)).astype(np.int64)
        to = rng.integers(0, 2, size=k).astype(np.int8)
        for which in (0, 1):  # one call per source cluster
            sel = rows[to == which]
            if len(sel) == 0:
                continue
            for e in (fast, slow):
                e.update_nodes(fgs.Packer(2).node_rows(snaps[which], sel), sel)
            src[sel] = which
        p0 = int(rng.integers(0, P - 5))
        small_t = fgs.Packer(2).pods(a.pending, p0, p0 + 1 + step % 4)
        for tab, lo, hi in ((small_t, p0, p0 + 1 + step % 4), (big, 0, P)):
            x, y = fast.evaluate(tab, want_planes=True), slow.evaluate(tab, want_planes=True)
            _same(x, y, N)
            exp = np.where(src[None, :] == 0, ra.status[lo:hi] == 0, rb.status[lo:hi] == 0)
            assert np.array_equal(x.feasible(N), exp)
    tm = fast.timings()
    assert tm["node_patches"] >= 14 and slow.timings()["node_patches"] == 0
    assert tm["node_patch_redos"] >= 1  # some of those rows did leave the analysis
    fast.close()
    slow.close()


def test_a_row_outside_the_node_analysis_is_caught_on_the_device(gpu_ok, oracle):
    """The patch kernel keeps the host's node analysis only while every patched row stays inside it. A bind that leaves a
    node's free cpu odd (every other node's is a multiple of 2^k) cannot be narrowed under the shifts in force: the next
    one-launch cycle declines on the device, the analysis is redone, and the answers are the oracle's."""
    S, N = 2, 900
    s = fgs.config_synth(2, n_nodes=N, n_pending=64)
    cfg = fgs.config_plugins(2)
    snap = L.Snapshot()
    C.memmove(C.byref(snap), s.snapshot, C.sizeof(L.Snapshot))
    req = np.ctypeslib.as_array(snap.requested, shape=(N, 3 + S)).copy()
    nz = np.ctypeslib.as_array(snap.nonzero, shape=(N, 2)).copy()
    gen = np.arange(1, N + 1, dtype=np.int64)
    snap.requested, snap.nonzero, snap.generation = (x.ctypes.data_as(L.p_i64) for x in (req, nz, gen))
    snap_p = C.pointer(snap)
    eng = _engine(cfg)
    pk = fgs.Packer(S)

    def cycle(p0, p1, planes):
        eng.sync_nodes(pk, snap_p)
        got = eng.evaluate(fgs.Packer(S).pods(s.pending, p0, p1), want_planes=planes)
        ref = oracle.evaluate(snap_p, s.pending, cfg, p0, p1, workers=2)
        assert np.array_equal(got.feasible_count, ref.feasible_count) and np.array_equal(got.pod_err, ref.pod_err)
        ok = (ref.pod_err == 0) & (ref.feasible_count >= 1)
        assert np.array_equal(got.selected_score[ok], ref.max_total[ok])
        idx = np.nonzero(ok)[0]
        assert np.array_equal(ref.total[idx, got.selected_node[idx]], ref.max_total[idx])
        if planes:
            assert np.array_equal(got.feasible(N), ref.status == 0)
            m = (ref.status == 0) & ok[:, None]
            assert np.array_equal(got.scores[:, :N][m], ref.total[m])
        return eng.timings()

    cycle(0, 2, False)
    g = N + 1
    BIG = 1 << 44  # NonZeroRequested memory of 16 TiB: above every node's maximum, and too large for the int32 kernels
    steps = [(500, 256 << 20, 0), (250, 1 << 20, 0), (0, 0, BIG), (1000, 1 << 30, 0), (0, 0, -BIG), (100, 1 << 20, 0), (200, 2 << 20, 0)]
    redos = 0
    for step, (d_cpu, d_mem, d_nz) in enumerate(steps):
        i = 17 if d_nz else 40 + 31 * step
        req[i, 0] += d_cpu
        nz[i, 0] += d_cpu
        req[i, 1] += d_mem
        nz[i, 1] += d_mem + d_nz
        gen[i] = g
        g += 1
        tm = cycle(step, step + 2, planes=bool(step % 2))
        assert tm["node_patches"] == step + 1
        if d_nz > 0:    # declined on the device, analysis redone; the batch then takes the int64 kernels
            assert tm["node_patch_redos"] == redos + 1 and not tm["small"], tm
        elif d_nz < 0:  # back inside; the loosened analysis is tightened once and the one-launch path returns
            assert tm["node_patch_redos"] == redos and tm["small"], tm
        elif step > 4:
            assert tm["small"], tm
        redos = tm["node_patch_redos"]
    # the general path settles a pending patch the same way (it waits for the flag before it reads the analysis)
    req[5, 1] += 3
    nz[5, 1] += 3
    gen[5] = g
    before = eng.timings()["node_patch_redos"]
    eng.sync_nodes(pk, snap_p)
    got = eng.evaluate(fgs.Packer(S).pods(s.pending), want_planes=True)
    ref = oracle.evaluate(snap_p, s.pending, cfg, workers=2)
    assert np.array_equal(got.feasible(N), ref.status == 0) and np.array_equal(got.feasible_count, ref.feasible_count)
    ok = (ref.pod_err == 0) & (ref.feasible_count >= 1)
    assert np.array_equal(got.selected_score[ok], ref.max_total[ok])
    assert eng.timings()["node_patch_redos"] in (before, before + 1)
    eng.close()


def test_a_node_listed_twice_takes_the_later_row(gpu_ok):
    a = fgs.config_synth(2, n_nodes=300, n_pending=50)
    b = fgs.config_synth(2, n_nodes=300, n_pending=50, seed_nodes=0x1234)
    cfg = fgs.config_plugins(2)
    t = fgs.Packer(2).pods(a.pending)
    for k in (3, 12):  # as kernel arguments, and through the compact buffer
        got, want = _engine(cfg), _engine(cfg, patch=False)
        for e in (got, want):
            e.upload_nodes(fgs.Packer(2).nodes(a.snapshot))
        rows = np.array(list(range(10, 10 + k)) + [10 + k - 1, 10], np.int64)       # table rows: b's nodes ...
        dirty = np.array(list(range(40, 40 + k)) + [40, 40 + k - 1], np.int64)      # ... onto these; 40 and 40+k-1 twice
        for e in (got, want):
            e.update_nodes(fgs.Packer(2).node_rows(b.snapshot, rows), dirty)
        _same(got.evaluate(t, want_planes=True), want.evaluate(t, want_planes=True), 300)
        got.close()
        want.close()


def test_a_batch_staged_before_a_node_change_is_refused(gpu_ok):
    """The staged batch's int32 columns follow the node table's shifts: passes over it after fgs_update_nodes are refused"""
    s = fgs.config_synth(2, n_nodes=500, n_pending=100)
    cfg = fgs.config_plugins(2)
    eng = _engine(cfg)
    pk = fgs.Packer(2)
    eng.upload_nodes(pk.nodes(s.snapshot))
    t = fgs.Packer(2).pods(s.pending)
    eng.stage_pods(t)
    eng.update_nodes(fgs.Packer(2).nodes(s.snapshot, 9, 10), [9])
    for call in (eng.pass_reduce, eng.pass_emit):
        with pytest.raises(fgs.FgsError) as ei:
            call()
        assert ei.value.code == L.FGS_E_STATE
    eng.stage_pods(t)
    eng.pass_reduce()
    eng.pass_emit()
    eng.fetch_results()
    eng.close()

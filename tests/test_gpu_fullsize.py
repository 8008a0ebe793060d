"""GPU tests at BASELINE.json's full sizes, where the oracle cannot cover every pair in seconds: size-independent
properties of the outputs (checked on the device with torch), agreement of the two kernel families on the whole
matrix, and the oracle on a sample of pods."""
import ctypes as C
import os

import numpy as np
import pytest

import flex_gpu_scheduler_b200 as fgs
from flex_gpu_scheduler_b200 import _lib as L

pytestmark = pytest.mark.gpu


class _Dev:
    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": typestr, "data": (int(ptr), False), "version": 2}


def _planes(torch, eng, P):
    v = eng.device_view()
    W, stride = int(v.words_per_row), int(v.score_row_stride)
    bm = torch.as_tensor(_Dev(v.feasible_bitmap, P * W, "<i4"), device="cuda").view(P, W)
    sc = torch.as_tensor(_Dev(v.scores, P * stride, "<i8"), device="cuda").view(P, stride)
    return bm, sc


def _run(force_int64, s, cfg):
    os.environ["FGS_FORCE_INT64"] = "1" if force_int64 else "0"
    os.environ["FGS_FILTER"] = "compare" if force_int64 else "lookup"  # the int64 family with its own Filter loop: an independent bitmap
    try:
        eng = fgs.Engine(device=0, n_scalars=2, total_nodes=s.n_nodes)
    finally:
        os.environ.pop("FGS_FORCE_INT64", None)
        os.environ.pop("FGS_FILTER", None)
    eng.set_plugins(cfg)
    pk = fgs.Packer(2)
    eng.upload_nodes(pk.nodes(s.snapshot))
    table = pk.pods(s.pending)
    res = eng.evaluate(table)
    return eng, table, res


def test_config2_full_size_properties(gpu_ok, oracle):
    import torch
    s = fgs.config_synth(2)  # 5 000 nodes x 100 000 pods
    cfg = fgs.config_plugins(2)
    N, P = s.n_nodes, s.n_pending
    eng, table, res = _run(False, s, cfg)
    t = eng.timings()
    assert t["reduce_lookup"] and t["emit_int32"]  # the benchmark configuration: Filter by table look-up, int32 emit kernel
    bm, sc = _planes(torch, eng, P)
    # feasibility as a [P][N] mask, in row chunks to bound memory
    shifts = torch.arange(32, device="cuda", dtype=torch.int32)
    counts = torch.zeros(P, dtype=torch.int64, device="cuda")
    best = torch.full((P,), -1, dtype=torch.int64, device="cuda")
    sel = torch.from_numpy(res.selected_node.astype(np.int64)).cuda()
    sel_ok = torch.ones(P, dtype=torch.bool, device="cuda")
    for p0 in range(0, P, 8192):
        p1 = min(P, p0 + 8192)
        bits = ((bm[p0:p1].unsqueeze(-1) >> shifts) & 1).view(p1 - p0, -1)[:, :N].bool()
        counts[p0:p1] = bits.sum(dim=1)
        masked = torch.where(bits, sc[p0:p1, :N], torch.full((), -1, dtype=torch.int64, device="cuda"))
        best[p0:p1] = masked.max(dim=1).values
        has = bits.any(dim=1)
        idx = sel[p0:p1].clamp(min=0)
        picked = bits.gather(1, idx.unsqueeze(1)).squeeze(1)
        sel_ok[p0:p1] = torch.where(has, picked & (sel[p0:p1] >= 0), sel[p0:p1] == -1)
    # 1. the per-pod count is the popcount of the pod's bitmap row
    assert torch.equal(counts.cpu(), torch.from_numpy(res.feasible_count.astype(np.int64)))
    # 2. selectHost picked a feasible node whose score is the row maximum over the feasible set
    assert bool(sel_ok.all())
    okpods = torch.from_numpy((res.pod_err == 0) & (res.feasible_count > 0)).cuda()
    assert torch.equal(best[okpods].cpu(), torch.from_numpy(res.selected_score)[okpods.cpu()])
    # 3. scores are weighted sums of two plugins normalised to [0,100]
    assert int(best.max()) <= 200
    checksum = int(sc[:, :N].sum())
    # 4. idempotence: evaluating the same batch again changes nothing
    res2 = eng.evaluate(table)
    assert np.array_equal(res2.feasible_count, res.feasible_count) and np.array_equal(res2.selected_node, res.selected_node)
    assert int(sc[:, :N].sum()) == checksum
    keep_sc = sc[:, :N].clone()
    keep_bm = bm.clone()
    eng.close()
    # 5. the int64 kernels produce the same bitmap, the same 5e8 scores and the same per-pod results
    eng64, table64, res64 = _run(True, s, cfg)
    t = eng64.timings()
    assert not t["reduce_int32"] and not t["emit_int32"]
    bm64, sc64 = _planes(torch, eng64, P)
    assert torch.equal(bm64, keep_bm)
    feas_any = torch.from_numpy(res.feasible_count > 0).cuda()
    # scores of infeasible pairs are unspecified: compare where the bitmap says feasible
    for p0 in range(0, P, 8192):
        p1 = min(P, p0 + 8192)
        bits = ((keep_bm[p0:p1].unsqueeze(-1) >> shifts) & 1).view(p1 - p0, -1)[:, :N].bool()
        assert torch.equal(sc64[p0:p1, :N][bits], keep_sc[p0:p1][bits])
    assert np.array_equal(res64.feasible_count, res.feasible_count) and np.array_equal(res64.pod_err, res.pod_err)
    assert np.array_equal(res64.selected_score[res.feasible_count > 0], res.selected_score[res.feasible_count > 0])
    eng64.close()
    # 6. the oracle on a sample of pods, all nodes
    sample = list(range(0, P, P // 48))[:48]
    for p in sample:
        ref = oracle.evaluate(s.snapshot, s.pending, cfg, p, p + 1, workers=4)
        f = ref.status[0] == 0
        assert int(f.sum()) == int(res.feasible_count[p])
        assert int(ref.pod_err[0]) == int(res.pod_err[p])
        if f.any() and ref.pod_err[0] == 0:
            assert np.array_equal(keep_sc[p].cpu().numpy()[f], ref.total[0][f])
            assert int(res.selected_score[p]) == int(ref.max_total[0])


def test_config3_and_4_full_size_against_oracle_sample(gpu_ok, oracle):
    # C3: 2 000 PodGroups over 10 000 nodes; C4: 500 namespaces, 100 000 pods (the quota check needs no nodes)
    s3 = fgs.config_synth(3)
    eng = fgs.Engine(device=0)
    eng.upload_nodes(fgs.Packer(2).nodes(s3.snapshot))
    pre, permit, assigned = eng.gang_check(s3.snapshot, s3.pending, s3.podgroups)
    lib = oracle.load()
    G = int(s3.podgroups.contents.n_groups)
    ref_assigned = np.zeros(G, np.int32)
    lib.fgs_oracle_gang_permit(s3.snapshot, s3.pending, s3.podgroups, 0, 0, None, ref_assigned.ctypes.data_as(L.p_i32))
    assert np.array_equal(assigned, ref_assigned)
    # CheckClusterResource through the oracle for every group that reaches the cluster walk is O(G·N·pods):
    # check a sample of groups exactly, through their first pending member
    pg = np.ctypeslib.as_array(s3.pending.contents.pg, shape=(s3.n_pending,))
    first = {}
    for i, g in enumerate(pg):
        if g >= 0 and g not in first:
            first[int(g)] = i
    groups = sorted(first)[::max(1, len(first) // 60)][:60]
    for g in groups:
        p = first[g]
        code = np.zeros(1, np.uint8)
        lib.fgs_oracle_gang_prefilter(s3.snapshot, s3.pending, s3.podgroups, p, p + 1, code.ctypes.data_as(L.p_u8))
        assert int(code[0]) == int(pre[p]), (g, p)
        st = np.zeros(1, np.uint8)
        lib.fgs_oracle_gang_permit(s3.snapshot, s3.pending, s3.podgroups, p, p + 1, st.ctypes.data_as(L.p_u8), None)
        assert int(st[0]) == int(permit[p]), (g, p)
    assert set(np.unique(pre)) >= {0, 2, 3}
    eng.close()
    s4 = fgs.config_synth(4, n_nodes=64)  # quota tables do not depend on the node count
    eng = fgs.Engine(device=0)
    code = eng.quota_check(fgs.Packer(2).pods(s4.pending), s4.quotas)
    ref = oracle.quota_prefilter(s4.pending, s4.quotas)
    assert np.array_equal(code, ref)
    eng.close()


# ---------------------------------------------------------------------------------------------------------------------
# BASELINE.json's configurations at their stated sizes against the oracle: C1 and C2 exhaustively — every
# (pod,node) pair —, C4 and C5 on hashed blocks of pods × ALL nodes. The planes stay on the device; the oracle's rows go
# up in chunks and are compared there.
# ---------------------------------------------------------------------------------------------------------------------

def _ncpu():
    return max(1, os.cpu_count() or 1)


def _compare_rows(torch, oracle, s, cfg, res, bm, sc, p0, p1, N, base=0, status=None):
    """pods [p0,p1) of the synthetic set = rows [p0-base, p1-base) of the engine's current batch: feasibility bits, count,
    error bits, totals on every feasible pair, selectHost's score, the chosen node inside the oracle's tie set, and (if
    given) the merged Filter status plane. Returns the number of (pod,node) pairs compared."""
    ref = oracle.evaluate(s.snapshot, s.pending, cfg, p0, p1, workers=_ncpu())
    r0, r1 = p0 - base, p1 - base
    shifts = torch.arange(32, device="cuda", dtype=torch.int32)
    bits = ((bm[r0:r1].unsqueeze(-1) >> shifts) & 1).view(r1 - r0, -1)[:, :N].bool()
    feas_ref = torch.from_numpy(ref.status == 0).cuda()
    assert torch.equal(bits, feas_ref), f"feasibility differs in pods [{p0},{p1})"
    assert np.array_equal(res.feasible_count[r0:r1], ref.feasible_count)
    assert np.array_equal(res.pod_err[r0:r1], ref.pod_err)
    if status is not None:
        assert np.array_equal(status[r0:r1], ref.status), f"status plane differs in pods [{p0},{p1})"
    ok = (ref.pod_err == 0) & (ref.feasible_count >= 1)  # scores of a cycle that errors out are not defined
    m = feas_ref & torch.from_numpy(ok).cuda()[:, None]
    tot_ref = torch.from_numpy(ref.total).cuda()
    assert torch.equal(sc[r0:r1, :N][m], tot_ref[m]), f"scores differ in pods [{p0},{p1})"
    idx = np.nonzero(ok)[0]
    sel = res.selected_node[r0:r1]
    assert np.all(sel[ref.feasible_count == 0] == -1)
    assert np.all(sel[idx] >= 0)
    assert np.array_equal(res.selected_score[r0:r1][idx], ref.max_total[idx])
    assert np.array_equal(ref.total[idx, sel[idx]], ref.max_total[idx]), "selected node outside the max-score tie set"
    assert np.all(ref.status[idx, sel[idx]] == 0)
    return (p1 - p0) * N


def _exhaustive(torch, oracle, cfg_id, chunk):
    s = fgs.config_synth(cfg_id)
    cfg = fgs.config_plugins(cfg_id)
    N, P = s.n_nodes, s.n_pending
    eng, table, res = _run(False, s, cfg)
    t = eng.timings()
    assert t["reduce_lookup"] and t["emit_int32"]
    bm, sc = _planes(torch, eng, P)
    status = eng.status_plane(0)
    done = 0
    for p0 in range(0, P, chunk):
        done += _compare_rows(torch, oracle, s, cfg, res, bm, sc, p0, min(P, p0 + chunk), N, status=status)
    assert done == N * P
    eng.close()
    return s, cfg


def test_config1_full_size_every_pair_against_the_oracle(gpu_ok, oracle):
    import torch
    _exhaustive(torch, oracle, 1, 2048)  # 1 000 nodes x 10 000 pods: 1e7 pairs


def test_config2_full_size_every_pair_against_the_oracle(gpu_ok, oracle):
    import torch
    _exhaustive(torch, oracle, 2, 2048)  # 5 000 nodes x 100 000 pods: 5e8 pairs


def _hashed_blocks(n_items, block, want, seed):
    """`want` distinct block indices of [0, n_items / block), chosen by a multiplicative hash (deterministic)"""
    n_blocks = n_items // block
    order = sorted(range(n_blocks), key=lambda b: ((b + seed) * 0x9E3779B1) & 0xffffffff)
    return sorted(order[:want])


def test_config4_full_size_quota_prefilter_and_fit_on_hashed_pods(gpu_ok, oracle):
    """C4 as BASELINE.json states it: 20 000 nodes, 500 namespaces, 100 000 pods — CapacityScheduling.PreFilter for every
    pod, and the FlexGPU + NodeResourcesFit evaluation on 2 048 hashed pods x all 20 000 nodes (4.1e7 pairs)"""
    import torch
    s = fgs.config_synth(4)
    cfg = fgs.config_plugins(4)
    N, P = s.n_nodes, s.n_pending
    assert (N, P) == (20000, 100000)
    eng, table, res = _run(False, s, cfg)
    code = eng.quota_check(table, s.quotas)
    ref_code = oracle.quota_prefilter(s.pending, s.quotas)
    assert np.array_equal(code, ref_code)
    assert set(np.unique(ref_code)) >= {0, 1}
    bm, sc = _planes(torch, eng, P)
    done = 0
    for b in _hashed_blocks(P, 16, 128, 4):
        done += _compare_rows(torch, oracle, s, cfg, res, bm, sc, 16 * b, 16 * b + 16, N)
    assert done == 2048 * N
    # the pods the quota admits are the ones whose evaluation the cycle goes on with: they are among the compared ones
    assert (ref_code[[16 * b for b in _hashed_blocks(P, 16, 128, 4)]] == 0).any()
    eng.close()


def test_config5_one_gpu_hashed_pods_all_nodes(gpu_ok, oracle):
    """C5 on ONE GPU: 50 000 nodes x 1 000 000 pods in 65 536-pod batches; 4 096 hashed pods (256 blocks of 16, spread over
    all batches) x all 50 000 nodes against the oracle: 2e8 pairs = 1/244 of the matrix, whole rows, so that the per-pod
    reductions (count, NormalizeScore maximum through the totals, selectHost) are covered too"""
    import torch
    s = fgs.config_synth(5)
    cfg = fgs.config_plugins(5)
    N, P = s.n_nodes, s.n_pending
    assert (N, P) == (50000, 1000000)
    eng = fgs.Engine(device=0, n_scalars=2, total_nodes=N)
    eng.set_plugins(cfg)
    eng.upload_nodes(fgs.Packer(2).nodes(s.snapshot))
    blocks = _hashed_blocks(P, 16, 256, 5)
    BATCH = 65536
    done = 0
    pk = fgs.Packer(2)
    for b0 in range(0, P, BATCH):
        b1 = min(P, b0 + BATCH)
        mine = [b for b in blocks if b0 <= 16 * b < b1]
        if not mine:
            continue
        res = eng.evaluate(pk.pods(s.pending, b0, b1))
        t = eng.timings()
        assert t["reduce_lookup"] and t["emit_int32"]
        bm, sc = _planes(torch, eng, b1 - b0)
        for b in mine:
            done += _compare_rows(torch, oracle, s, cfg, res, bm, sc, 16 * b, 16 * b + 16, N, base=b0)
    assert done == 4096 * N
    eng.close()

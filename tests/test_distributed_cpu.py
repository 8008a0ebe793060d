"""CPU tests of the N>1 path: two processes over gloo play two node shards. The kernels cannot run here, so
each rank takes its shard's per-pod partial results from the oracle evaluated on ITS node range only (counts,
raw flex maxima, packed candidate keys), pushes them through the same merge functions bench.py uses on the
GPUs (flex_gpu_scheduler_b200.distributed) and checks the merged vectors against the oracle on the whole cluster."""
import ctypes as C
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _sub_snapshot(L, snap, n0, n1):
    """the shard [n0,n1) as its own fgs_snapshot: per-node arrays advanced, pod_off left absolute"""
    S = snap.n_scalars
    sub = L.Snapshot()
    C.memmove(C.byref(sub), C.byref(snap), C.sizeof(sub))
    sub.n_nodes = n1 - n0

    def adv(p, ctype, k):
        return C.cast(C.addressof(p.contents) + k * C.sizeof(ctype), C.POINTER(ctype))

    sub.node_valid = adv(snap.node_valid, C.c_uint8, n0)
    sub.alloc = adv(snap.alloc, C.c_int64, n0 * (4 + S))
    sub.alloc_scalar_present = adv(snap.alloc_scalar_present, C.c_uint32, n0)
    sub.requested = adv(snap.requested, C.c_int64, n0 * (3 + S))
    sub.nonzero = adv(snap.nonzero, C.c_int64, n0 * 2)
    sub.pod_off = adv(snap.pod_off, C.c_int64, n0)
    return sub


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    try:
        import torch
        import torch.distributed as dist
        import flex_gpu_scheduler_b200 as fgs
        from flex_gpu_scheduler_b200 import _lib as L, distributed as D
        from oracle import oracle
        dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
        N, P = 333, 120
        s = fgs.config_synth(2, n_nodes=N, n_pending=P)
        # filter + raw FlexGPU score only: quantities that do not depend on a cross-shard normaliser
        cfg = fgs.plugin_cfg(filter_plugins=("NodeResourcesFit", "FlexGPU"), score_plugins=(("FlexGPU", 1),))
        n0, n1 = D.shard_bounds(N, world, rank)
        sub = _sub_snapshot(L, s.snapshot.contents, n0, n1)
        part = oracle.evaluate(C.pointer(sub), s.pending, cfg, workers=1, per_plugin=True)
        feas = part.status == 0
        raw = np.where(feas, part.raw_flex, np.iinfo(np.int64).min)
        p_pad = (P + 63) // 64 * 64
        rm = torch.full((D.REDUCE_ROWS * p_pad,), np.iinfo(np.int64).min, dtype=torch.int64)
        rm[:P] = torch.from_numpy(raw.max(axis=1) if n1 > n0 else np.full(P, np.iinfo(np.int64).min))
        rs = torch.zeros(p_pad, dtype=torch.int64)
        rs[:P] = torch.from_numpy(part.feasible_count.astype(np.int64))
        D.merge_after_reduce(dist, rm, rs)
        # candidate keys in the engine's layout: (raw score << 40 | global node) + 1, best node of the shard
        sk = torch.zeros(D.KEY_ROWS * p_pad, dtype=torch.int64)
        best = np.where(feas, part.raw_flex, -1)
        arg = best.argmax(axis=1)
        has = feas.any(axis=1)
        key = ((best[np.arange(P), arg].astype(np.int64) << 40) | (arg + n0)) + 1
        sk[:P] = torch.from_numpy(np.where(has, key, 0))
        D.merge_after_emit(dist, sk)
        if rank == 0:
            whole = oracle.evaluate(s.snapshot, s.pending, cfg, workers=1, per_plugin=True)
            wf = whole.status == 0
            assert np.array_equal(rs[:P].numpy(), whole.feasible_count)
            want_max = np.where(wf, whole.raw_flex, np.iinfo(np.int64).min).max(axis=1)
            assert np.array_equal(rm[:P].numpy(), want_max)
            node, score = D.decode_keys(sk[:P].numpy())
            anyf = wf.any(axis=1)
            assert np.all(node[~anyf] == -1)
            idx = np.nonzero(anyf)[0]
            assert np.all(wf[idx, node[idx]])
            assert np.array_equal(whole.raw_flex[idx, node[idx]], np.where(wf, whole.raw_flex, -1).max(axis=1)[idx])
            assert np.array_equal(score[idx], whole.raw_flex[idx, node[idx]])
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        import traceback
        q.put((rank, "FAIL: " + traceback.format_exc()))


def test_two_shards_merge_like_one(oracle):
    import multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(r[1] == "ok" for r in res), res


def test_shard_bounds_cover_without_overlap():
    sys.path.insert(0, ROOT)
    from flex_gpu_scheduler_b200 import distributed as D
    for n in (0, 1, 7, 5000, 50001):
        for w in (1, 2, 4, 8):
            b = [D.shard_bounds(n, w, r) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            assert all(0 <= hi - lo <= (n + w - 1) // w for lo, hi in b)


def test_decode_keys():
    sys.path.insert(0, ROOT)
    from flex_gpu_scheduler_b200 import distributed as D
    keys = np.array([0, ((150 << 40) | (0xBEEF << 24) | 4999) + 1, ((0 << 40) | 0) + 1], np.int64)
    node, score = D.decode_keys(keys)
    assert node.tolist() == [-1, 4999, 0] and score.tolist() == [0, 150, 0]

"""Run under torchrun with N >= 2 GPUs: node-sharded evaluation against the oracle on the whole cluster, with the
merges as NCCL all-reduces, as the engine's peer-memory kernels through NVSwitch multicast, and as the same kernels
over plain peer pointers. Launched by tests/test_gpu_multi.py; exits non-zero on any mismatch."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def check_calls_with_merges(fgs, eng, s, ref, P):
    """fgs_evaluate and the submit / collect pipeline of a sharded engine include the merges (every rank makes the
    same calls): per-pod results against the oracle's, with batches of different sizes alternating in the two slots"""
    def same(r, b0, b1):
        ok = (ref.pod_err[b0:b1] == 0) & (ref.feasible_count[b0:b1] >= 1)
        assert np.array_equal(r.feasible_count, ref.feasible_count[b0:b1]), "counts"
        assert np.array_equal(r.pod_err, ref.pod_err[b0:b1]), "error bits"
        assert np.array_equal(r.selected_score[ok], ref.max_total[b0:b1][ok]), "selected score"
        idx = np.nonzero(ok)[0]
        feas = ref.status[b0:b1] == 0
        assert np.all(feas[idx, r.selected_node[idx]]), "selected node infeasible"
        assert np.array_equal(ref.total[b0:b1][idx, r.selected_node[idx]], ref.max_total[b0:b1][idx]), "selected node outside tie set"
        assert np.all(r.selected_node[ref.feasible_count[b0:b1] == 0] == -1)

    h = P // 3
    cuts = [(0, h), (h, P), (0, P)]
    packers = [fgs.Packer(2) for _ in cuts]
    tabs = [pk.pods(s.pending, b0, b1) for pk, (b0, b1) in zip(packers, cuts)]
    same(eng.evaluate(tabs[2]), 0, P)
    same(eng.evaluate(tabs[0]), 0, h)
    order = [0, 1, 2, 1, 0, 0, 2]
    for depth in (1, 2, 3):  # batches in flight (a sharded engine takes up to three)
        q = []
        for i in order:
            eng.submit(tabs[i])
            q.append(i)
            if len(q) == depth:
                same(eng.collect(), *cuts[q.pop(0)])
        while q:
            same(eng.collect(), *cuts[q.pop(0)])
    # the three batches left resident in the ring of slots again, without staging: the ring is walked round-robin, so
    # the slot after the last one filled comes first — it holds the third-to-last batch
    ring = [cuts[order[-3]], cuts[order[-2]], cuts[order[-1]]]
    for c in ring:
        eng.resubmit(c[1] - c[0])
    same(eng.collect(), *ring[0])
    eng.resubmit(ring[0][1] - ring[0][0])
    for c in ring[1:] + ring[:1]:
        same(eng.collect(), *c)
    eng.peer_status()


def check_growing_batches(fgs, eng, s, ref, P):
    """a batch larger than any before makes the engine replace its bitmap / score planes while the previous batch still
    waits for its emit pass (the sharded pipeline defers it by one submit): results must not change"""
    sizes = [min(P, 300), min(P, 700), P]
    packers = [fgs.Packer(2, pin=(k % 2 == 1)) for k in range(len(sizes))]
    tabs = [pk.pods(s.pending, 0, n) for pk, n in zip(packers, sizes)]
    for t in tabs:
        eng.submit(t)
    for n in sizes:
        r = eng.collect()
        ok = (ref.pod_err[:n] == 0) & (ref.feasible_count[:n] >= 1)
        assert np.array_equal(r.feasible_count, ref.feasible_count[:n]), "growing batches: counts"
        assert np.array_equal(r.pod_err, ref.pod_err[:n]), "growing batches: error bits"
        assert np.array_equal(r.selected_score[ok], ref.max_total[:n][ok]), "growing batches: selected score"
    eng.peer_status()


def check_masked_scoring(fgs, eng, s, cfg, table, oracle, N, P, n0, n1):
    """fgs_score_masked on a sharded engine: every rank passes its shard's columns of the same feasible mask"""
    rng = np.random.default_rng(1234 + N)
    keep = rng.uniform(0.3, 0.9, size=(P, 1))
    mask = rng.random((P, N)) < keep
    mask[P // 2] = False
    eng.evaluate(table, skip_pass_b=1)
    got = eng.score_masked(mask[:, n0:n1], want_planes=True)
    ref = oracle.evaluate(s.snapshot, s.pending, cfg, workers=2, ext_mask=mask.astype(np.uint8))
    feas = ref.status == 0
    assert np.array_equal(got.feasible(n1 - n0), feas[:, n0:n1]), "masked: bitmap shard"
    assert np.array_equal(got.feasible_count, ref.feasible_count), "masked: merged counts"
    assert np.array_equal(got.pod_err, ref.pod_err), "masked: error bits"
    ok = (ref.pod_err == 0) & (ref.feasible_count >= 1)
    m = feas[:, n0:n1] & ok[:, None]
    assert np.array_equal(got.scores[:, :n1 - n0][m], ref.total[:, n0:n1][m]), "masked: scores (normalised over the masked list of ALL shards)"
    idx = np.nonzero(ok)[0]
    assert np.array_equal(got.selected_score[idx], ref.max_total[idx]), "masked: selected score"
    assert np.array_equal(ref.total[idx, got.selected_node[idx]], ref.max_total[idx]) and np.all(feas[idx, got.selected_node[idx]])
    assert np.all(got.selected_node[ref.feasible_count == 0] == -1)


def check_mixed_shift_batches(fgs, D, oracle, torch, dist, world, rank, local):
    """batches that need different int32 shifts, in flight together on a sharded engine: the emit pass of a batch is issued
    by the NEXT submit, after that batch was staged — it must still read the node table narrowed for its own shifts"""
    import builders as B
    import mixed_shifts as M
    N = 2600
    snap = B.build_snapshot(M.cluster(N))
    cfg = fgs.config_plugins(2)
    sets = {k: B.build_podset(M.batch(k, n, seed)) for k, n, seed in (("A", 900, 1), ("B", 700, 2), ("C", 1100, 3))}
    refs = {k: oracle.evaluate(snap.ptr, v.ptr, cfg, workers=2) for k, v in sets.items()}
    n0, n1 = D.shard_bounds(N, world, rank)
    stream = torch.cuda.Stream()
    eng = fgs.Engine(device=local, n_scalars=2, node_index_base=n0, total_nodes=N, stream=stream.cuda_stream)
    eng.set_plugins(cfg)
    eng.upload_nodes(fgs.Packer(2).nodes(snap.ptr, n0, n1))
    px, why = D.try_peer_exchange(torch, dist, eng, 1100, multicast=True)
    if px is None:
        if rank == 0:
            print("mixed-shift batches: no peer exchange:", why)
        eng.close()
        return
    packers = {k: fgs.Packer(2, pin=(k != "B")) for k in sets}
    tabs = {k: packers[k].pods(v.ptr) for k, v in sets.items()}

    def same(got, ref):
        assert np.array_equal(got.feasible_count, ref.feasible_count), "mixed shifts: counts"
        assert np.array_equal(got.pod_err, ref.pod_err), "mixed shifts: error bits"
        ok = (ref.pod_err == 0) & (ref.feasible_count >= 1)
        assert np.array_equal(got.selected_score[ok], ref.max_total[ok]), "mixed shifts: selected score"
        idx = np.nonzero(ok)[0]
        assert np.array_equal(ref.total[idx, got.selected_node[idx]], ref.max_total[idx]), "mixed shifts: selected node outside tie set"

    depth = eng.max_in_flight()
    for order in ("ABCABC", "AABBCC", "CBACAB"):
        q = []
        for k in order:
            eng.submit(tabs[k])
            q.append(k)
            if len(q) == depth:
                same(eng.collect(), refs[q.pop(0)])
        while q:
            same(eng.collect(), refs[q.pop(0)])
    # node rows change on every shard (row patches on the device), then the same batches again
    nodes = M.cluster(N)
    spare = M.cluster(N, seed=77)
    rnd = np.random.default_rng(5)
    for rounds in range(3):
        changed = np.sort(rnd.choice(N, size=[1, 6, 40][rounds], replace=False))
        for i in changed:
            nodes[i], spare[i] = spare[i], nodes[i]
        snap2 = B.build_snapshot(nodes)
        mine = changed[(changed >= n0) & (changed < n1)].astype(np.int64)
        if len(mine):
            eng.update_nodes(fgs.Packer(2).node_rows(snap2.ptr, mine), mine - n0)
        refs2 = {k: oracle.evaluate(snap2.ptr, v.ptr, cfg, workers=2) for k, v in sets.items()}
        q = []
        for k in "BACB":
            eng.submit(tabs[k])
            q.append(k)
            if len(q) == depth:
                same(eng.collect(), refs2[q.pop(0)])
        while q:
            same(eng.collect(), refs2[q.pop(0)])
        same(eng.evaluate(tabs["A"]), refs2["A"])
    eng.peer_status()
    px.close()
    eng.close()


def main():
    import torch
    import torch.distributed as dist
    import flex_gpu_scheduler_b200 as fgs
    from flex_gpu_scheduler_b200 import distributed as D
    from oracle import oracle

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    modes_run = []
    for mode, (cfg_id, N, P) in [(m, c) for m in ("nccl", "multicast", "peer") for c in ((2, 2311, 700), (1, 1500, 300), (2, 900, 4001))]:
        s = fgs.config_synth(cfg_id, n_nodes=N, n_pending=P, pm_quirk_nodes=80)
        cfg = fgs.config_plugins(cfg_id)
        n0, n1 = D.shard_bounds(N, world, rank)
        stream = torch.cuda.Stream()
        eng = fgs.Engine(device=local, n_scalars=2, node_index_base=n0, total_nodes=N, stream=stream.cuda_stream)
        eng.set_plugins(cfg)
        eng.upload_nodes(fgs.Packer(2).nodes(s.snapshot, n0, n1))
        ref = oracle.evaluate(s.snapshot, s.pending, cfg, workers=2)
        px = None
        if mode != "nccl":
            px, why = D.try_peer_exchange(torch, dist, eng, P, multicast=(mode == "multicast"))
            if px is None or (mode == "multicast" and not px.multicast):
                if rank == 0:
                    print("mode", mode, "unavailable:", why or "no multicast address")
                if px is not None:
                    px.close()
                eng.close()
                continue
        if px is not None:
            check_growing_batches(fgs, eng, s, ref, P)
            table = fgs.Packer(2).pods(s.pending)
            for _ in range(3):  # repeated merges: epochs and buffer reuse
                res = D.evaluate_sharded(torch, dist, eng, table, peer=px)
            assert eng.peer_status() == (mode == "multicast")
            eng.stage_pods(table)
            eng.pass_reduce()
            eng.peer_merge_reduce()
            eng.pass_emit()
            eng.peer_merge_select()
            res = eng.fetch_results(want_planes=True)
            eng.peer_status()
        else:
          with torch.cuda.stream(stream):
              table = fgs.Packer(2).pods(s.pending)
              eng.stage_pods(table)
              rm, rs, sk = D.device_views(torch, eng.device_view())
              eng.pass_reduce()
              # config 2 does not score NodeResourcesAllocatable: the count all-reduce may overlap the emit pass
              pending = D.merge_after_reduce(dist, rm, rs, defer_counts=(cfg_id == 2))
              eng.pass_emit()
              D.merge_after_emit(dist, sk)
              if pending is not None:
                  pending.wait()
              res = eng.fetch_results(want_planes=True)
        if px is not None:
            check_calls_with_merges(fgs, eng, s, ref, P)
            check_masked_scoring(fgs, eng, s, cfg, table, oracle, N, P, n0, n1)
        feas_ref = ref.status == 0
        # this shard's planes against the oracle's columns
        assert np.array_equal(res.feasible(n1 - n0), feas_ref[:, n0:n1]), "bitmap shard"
        ok = (ref.pod_err == 0) & (ref.feasible_count >= 1)
        m = feas_ref[:, n0:n1] & ok[:, None]
        assert np.array_equal(res.scores[:, :n1 - n0][m], ref.total[:, n0:n1][m]), "score shard (normalised over ALL shards)"
        # merged per-pod results are identical on every rank
        assert np.array_equal(res.feasible_count, ref.feasible_count), "merged counts"
        assert np.array_equal(res.pod_err, ref.pod_err), "merged error bits"
        idx = np.nonzero(ok)[0]
        sel = res.selected_node
        assert np.all(sel[ref.feasible_count == 0] == -1)
        assert np.all(feas_ref[idx, sel[idx]]) and np.array_equal(ref.total[idx, sel[idx]], ref.max_total[idx]), "selected node outside tie set"
        assert np.array_equal(res.selected_score[idx], ref.max_total[idx])
        if px is not None:
            px.close()
        eng.close()
        modes_run.append(mode)
    check_mixed_shift_batches(fgs, D, oracle, torch, dist, world, rank, local)
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("multi_gpu_parity ok, world", world, "modes", sorted(set(modes_run)))


if __name__ == "__main__":
    main()

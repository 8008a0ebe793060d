"""tools/merge_micro.py — the peer-memory merge kernels alone (back to back, no evaluation between them): per-call time
of fgs_peer_merge_reduce / fgs_peer_merge_select per mode and grid size. Run under torchrun with N >= 2 GPUs."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    import flex_gpu_scheduler_b200 as fgs
    from flex_gpu_scheduler_b200 import distributed as D

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    P, N = 100000, 1024 * world
    s = fgs.config_synth(2, n_nodes=N, n_pending=P)
    n0, n1 = D.shard_bounds(N, world, rank)
    stream = torch.cuda.Stream()
    rows = []
    for mode in ("multicast", "peer", "nccl"):
        for ctas in ((32, 64, 128) if mode != "nccl" else (0,)):
            os.environ["FGS_PEER_CTAS"] = str(ctas)
            eng = fgs.Engine(device=local, n_scalars=2, node_index_base=n0, total_nodes=N, stream=stream.cuda_stream)
            eng.set_plugins(fgs.config_plugins(2))
            eng.upload_nodes(fgs.Packer(2).nodes(s.snapshot, n0, n1))
            px = None
            if mode != "nccl":
                px, why = D.try_peer_exchange(torch, dist, eng, P, multicast=(mode == "multicast"))
                if px is None or (mode == "multicast" and not px.multicast):
                    eng.close()
                    continue
            with torch.cuda.stream(stream):
                eng.stage_pods(fgs.Packer(2).pods(s.pending))
                eng.pass_reduce()
                eng.pass_emit()
                rm, rs, sk = D.device_views(torch, eng.device_view())
                for which in ("reduce", "select"):
                    def call():
                        if px is not None:
                            (eng.peer_merge_reduce if which == "reduce" else eng.peer_merge_select)()
                        elif which == "reduce":
                            dist.all_reduce(rm, op=dist.ReduceOp.MAX)
                        else:
                            dist.all_reduce(sk, op=dist.ReduceOp.MAX)
                    for _ in range(5):
                        call()
                    torch.cuda.synchronize()
                    dist.barrier()
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record(stream)
                    for _ in range(50):
                        call()
                    b.record(stream)
                    torch.cuda.synchronize()
                    t = torch.tensor([a.elapsed_time(b) / 50], device="cuda", dtype=torch.float64)
                    dist.all_reduce(t, op=dist.ReduceOp.MAX)
                    rows.append((mode, ctas, which, float(t.item()) * 1e3))
                if px is not None:
                    eng.peer_status()
            if px is not None:
                px.close()
            eng.close()
    if rank == 0:
        for r in rows:
            print("world %d  %-9s ctas %3d  merge_%-6s %7.1f us" % ((world,) + r))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

"""Node-sharded evaluation over several GPUs: one process (and one libfgs engine) per GPU.

The path shards by NODES: rank r holds the contiguous snapshot range shard_bounds(N, world, r),
pods are replicated. Only per-pod vectors cross NVLink:
    after the reduce pass   reduce_max  int64 [6][p_pad]  all_reduce(MAX)   flex max, Allocatable hi / −lo, flags
                            reduce_sum  int64 [p_pad]     all_reduce(SUM)   feasible counts
    after the emit pass     select_key  int64 [3][p_pad]  all_reduce(MAX)   packed (score, hash, node) key + flags
On GPUs the merges are the engine's own kernels over NVLink peer memory (fgs_peer_merge_reduce / _select,
csrc/peer_merge.cu): PeerExchange only does the plumbing — an exchange buffer per rank in torch symmetric memory,
mapped by every rank, with the NVSwitch multicast address where there is one. Without symmetric memory (or in the
gloo CPU tests of the host logic) the same merges run as torch.distributed all-reduces on zero-copy views of the
engine's device buffers. The score matrix and the bitmap never leave their shard.
"""
import numpy as np

REDUCE_ROWS = 6
KEY_ROWS = 3


def shard_bounds(n_total, world, rank):
    """contiguous, balanced node range [n0, n1) of `rank`"""
    per = (n_total + world - 1) // world
    n0 = min(n_total, rank * per)
    return n0, min(n_total, n0 + per)


class _CudaArray:
    """__cuda_array_interface__ holder: lets torch wrap a device pointer the engine owns without copying"""

    def __init__(self, ptr, n_elems):
        self.__cuda_array_interface__ = {"shape": (int(n_elems),), "typestr": "<i8", "data": (int(ptr), False), "version": 2}


def device_views(torch, view):
    """(reduce_max, reduce_sum, select_key) as int64 torch tensors over the engine's buffers"""
    p = int(view.p_pad)
    return (torch.as_tensor(_CudaArray(view.reduce_max, REDUCE_ROWS * p), device="cuda"),
            torch.as_tensor(_CudaArray(view.reduce_sum, p), device="cuda"),
            torch.as_tensor(_CudaArray(view.select_key, KEY_ROWS * p), device="cuda"))


class PeerExchange:
    """Exchange buffers for the engine's peer-memory merges. Collective: every rank constructs it at the same point.
    Raises when symmetric memory cannot be set up (callers then stay on the all-reduce path)."""

    def __init__(self, torch, dist, engine, max_pods, multicast=True):
        import torch.distributed._symmetric_memory as symm
        self.engine = engine
        nbytes = engine.peer_bytes(max_pods)
        dev = torch.device("cuda", torch.cuda.current_device())
        self.buf = symm.empty(nbytes, dtype=torch.uint8, device=dev)
        self.hdl = symm.rendezvous(self.buf, dist.group.WORLD)
        rank, world = int(self.hdl.rank), int(self.hdl.world_size)
        ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        off = int(self.buf.data_ptr()) - ptrs[rank]  # the tensor's place inside the symmetric allocation
        mc = int(self.hdl.multicast_ptr) if multicast else 0
        engine.peer_attach(rank, world, [p + off for p in ptrs], mc + off if mc else 0, nbytes)
        self.multicast = bool(mc)
        torch.cuda.synchronize()
        dist.barrier()  # every header is zero before anyone merges

    def close(self):
        self.engine.peer_detach()
        self.buf = self.hdl = None


def try_peer_exchange(torch, dist, engine, max_pods, multicast=True):
    """(PeerExchange, None), or (None, why) when this box / torch build cannot do it; agreed across ranks"""
    px, why = None, None
    try:
        px = PeerExchange(torch, dist, engine, max_pods, multicast)
    except Exception as ex:  # noqa: BLE001 — whatever the rendezvous raises, the all-reduce path still works
        why = "%s: %s" % (type(ex).__name__, str(ex).splitlines()[0][:200] if str(ex) else "")
    flag = torch.tensor([1 if px is not None else 0], device="cuda", dtype=torch.int32)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if int(flag.item()) == 0:
        if px is not None:
            px.close()
        return None, why or "a peer rank could not set up symmetric memory"
    return px, None


def merge_after_reduce(dist, reduce_max, reduce_sum, defer_counts=False):
    """cross-shard merge between the two passes; tensors are modified in place.
    defer_counts: the emit pass needs the merged maxima but, unless NodeResourcesAllocatable is scored, not the
    merged feasible counts (only the per-pod epilogue does): their all-reduce is then issued asynchronously AFTER the
    MAX one, overlaps the emit pass, and the returned handle must be wait()ed before fgs_fetch_results."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return None
    dist.all_reduce(reduce_max, op=dist.ReduceOp.MAX)
    if defer_counts:
        return dist.all_reduce(reduce_sum, op=dist.ReduceOp.SUM, async_op=True)
    dist.all_reduce(reduce_sum, op=dist.ReduceOp.SUM)
    return None


def merge_after_emit(dist, select_key):
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return
    dist.all_reduce(select_key, op=dist.ReduceOp.MAX)


def decode_keys(keys, key_shift=40):
    """packed selectHost keys → (global node index or -1, total score); the layout of finalize_pods"""
    keys = np.asarray(keys, dtype=np.int64)
    k = keys.astype(np.uint64) - np.uint64(1)
    node = np.where(keys > 0, (k & np.uint64(0xFFFFFF)).astype(np.int64), -1)
    score = np.where(keys > 0, (k >> np.uint64(key_shift)).astype(np.int64), 0)
    return node, score


def evaluate_sharded(torch, dist, engine, pod_table, defer_counts=False, peer=None):
    """One full evaluation of `pod_table` on this rank's shard with the cross-shard merges in place.
    peer: a PeerExchange — the merges are then the engine's own kernels on its stream. Otherwise all-reduces, and this
    must run with the engine's stream current (torch.cuda.stream(...)) so that the collectives are ordered.
    defer_counts: see merge_after_reduce (only valid when NodeResourcesAllocatable is not a score plugin)."""
    engine.stage_pods(pod_table)
    if peer is not None:
        engine.pass_reduce()
        engine.peer_merge_reduce()
        engine.pass_emit()
        engine.peer_merge_select()
        return engine.fetch_results()
    view = engine.device_view()
    rm, rs, sk = device_views(torch, view)
    engine.pass_reduce()
    pending = merge_after_reduce(dist, rm, rs, defer_counts)
    engine.pass_emit()
    merge_after_emit(dist, sk)
    if pending is not None:
        pending.wait()
    return engine.fetch_results()

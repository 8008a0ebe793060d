"""Node-sharded evaluation over several GPUs: one process (and one libfgs engine) per GPU.

The path shards by NODES: rank r holds the contiguous snapshot range shard_bounds(N, world, r),
pods are replicated. Only per-pod vectors cross NVLink:
    after the reduce pass   reduce_max  int64 [6][p_pad]  all_reduce(MAX)   flex max, Allocatable hi / −lo, flags
                            reduce_sum  int64 [p_pad]     all_reduce(SUM)   feasible counts
    after the emit pass     select_key  int64 [3][p_pad]  all_reduce(MAX)   packed (score, hash, node) key + flags
The collectives run through torch.distributed (NCCL on GPUs, gloo in the CPU tests) on zero-copy views of the
engine's device buffers; the score matrix and the bitmap never leave their shard.
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


def evaluate_sharded(torch, dist, engine, pod_table, defer_counts=False):
    """One full evaluation of `pod_table` on this rank's shard with the cross-shard merges in place.
    Must run with the engine's stream current (torch.cuda.stream(...)) so that the collectives are ordered.
    defer_counts: see merge_after_reduce (only valid when NodeResourcesAllocatable is not a score plugin)."""
    engine.stage_pods(pod_table)
    view = engine.device_view()
    rm, rs, sk = device_views(torch, view)
    engine.pass_reduce()
    pending = merge_after_reduce(dist, rm, rs, defer_counts)
    engine.pass_emit()
    merge_after_emit(dist, sk)
    if pending is not None:
        pending.wait()
    return engine.fetch_results()

"""Would the reduce pass of one batch overlap the emit pass of another? Two engines on one GPU, each on its own stream,
issued so that A's reduce pass meets B's emit pass; compared with the same work on one stream. python tools/overlap_probe.py"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import flex_gpu_scheduler_b200 as fgs  # noqa: E402

s = fgs.config_synth(2)
engs = []
for _ in range(2):
    e = fgs.Engine(device=0)
    e.set_plugins(fgs.config_plugins(2))
    e.upload_nodes(fgs.Packer(2).nodes(s.snapshot))
    e.stage_pods(fgs.Packer(2).pods(s.pending))
    e.pass_reduce(); e.pass_emit(); e.synchronize()
    engs.append(e)
a, b = engs
K = 20


def sync():
    a.synchronize(); b.synchronize()


sync()
t0 = time.perf_counter()
for _ in range(K):
    a.pass_reduce(); a.pass_emit()
a.synchronize()
t1 = time.perf_counter()
print("one stream          : %.3f ms per batch" % (1e3 * (t1 - t0) / K))
sync()
b.pass_reduce()
t0 = time.perf_counter()
for _ in range(K):  # per iteration: two batches (one per engine); A's reduce is issued while B's emit is queued, and so on
    a.pass_reduce(); b.pass_emit()
    b.pass_reduce(); a.pass_emit()
sync()
t1 = time.perf_counter()
print("two streams, skewed : %.3f ms per batch" % (1e3 * (t1 - t0) / (2 * K)))

"""Host-side time of fgs_submit / fgs_collect in a steady stream of C2 batches, with two and with three batches in
flight (FGS_PIPE_SLOTS): python tools/pipe_breakdown.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import flex_gpu_scheduler_b200 as fgs

s = fgs.config_synth(2)
for depth in (2, 3):
    os.environ["FGS_PIPE_SLOTS"] = str(depth)
    eng = fgs.Engine(device=0)
    eng.set_plugins(fgs.config_plugins(2))
    eng.upload_nodes(fgs.Packer(2).nodes(s.snapshot))
    t = fgs.Packer(2).pods(s.pending)
    for _ in range(depth):
        eng.submit(t)
    for _ in range(depth):
        eng.collect()
    K = 40
    ts, tc, n = 0.0, 0.0, 0
    t0 = time.perf_counter()
    inflight = 0
    for _ in range(K):
        a = time.perf_counter(); eng.submit(t); b = time.perf_counter()
        ts += b - a
        inflight += 1
        if inflight == depth:
            a = time.perf_counter(); r = eng.collect(); b = time.perf_counter()
            tc += b - a; n += 1
            inflight -= 1
    for _ in range(inflight):
        r = eng.collect()
    wall = time.perf_counter() - t0
    print("depth %d  per step: wall %.3f ms, submit %.3f ms, collect (wait + unpack) %.3f ms" % (depth, 1e3 * wall / K, 1e3 * ts / K, 1e3 * tc / max(n, 1)))
    eng.close()

"""Host-side time of fgs_submit / fgs_collect in a steady stream of C2 batches: python tools/pipe_breakdown.py
(one B200: wall 1.26 ms per step = submit 0.27 ms + collect 0.98 ms, i.e. the host waits for the GPU)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import flex_gpu_scheduler_b200 as fgs

s = fgs.config_synth(2)
eng = fgs.Engine(device=0)
eng.set_plugins(fgs.config_plugins(2))
eng.upload_nodes(fgs.Packer(2).nodes(s.snapshot))
t = fgs.Packer(2).pods(s.pending)
eng.submit(t); eng.submit(t); eng.collect(); eng.collect()
K = 40
ts, tc = 0.0, 0.0
t0 = time.perf_counter()
eng.submit(t)
for _ in range(K - 1):
    a = time.perf_counter(); eng.submit(t); b = time.perf_counter(); r = eng.collect(); c = time.perf_counter()
    ts += b - a; tc += c - b
r = eng.collect()
wall = time.perf_counter() - t0
print("per step: wall %.3f ms, submit %.3f ms, collect (wait + unpack) %.3f ms" % (1e3 * wall / K, 1e3 * ts / (K - 1), 1e3 * tc / (K - 1)))

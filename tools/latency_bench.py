"""Latency of the calls a scheduler makes for ONE pod (or a handful), 5 000-node C2 cluster, one B200:
fgs_evaluate for 1 / 16 / 256 / 4096 pods, fgs_select_victims (+ pick) and fgs_fit_error_histogram.
python tools/latency_bench.py"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import flex_gpu_scheduler_b200 as fgs  # noqa: E402

s = fgs.config_synth(2)
eng = fgs.Engine(device=0)
eng.set_plugins(fgs.config_plugins(2))
eng.upload_nodes(fgs.Packer(2).nodes(s.snapshot))


def timed(f, reps=200):
    for _ in range(5):
        f()
    t0 = time.perf_counter()
    for _ in range(reps):
        f()
    return 1e6 * (time.perf_counter() - t0) / reps


out = {}
for P in (1, 16, 256, 4096):
    pk = fgs.Packer(2)
    t = pk.pods(s.pending, 0, P)
    out["evaluate_%d_pods_us" % P] = round(timed(lambda: eng.evaluate(t)), 1)
t1 = fgs.Packer(2).pods(s.pending, 0, 64)
eng.evaluate(t1)
out["fit_error_histogram_64_pods_us"] = round(timed(lambda: eng.fit_error_histogram(), 50), 1)
out["select_victims_and_pick_us"] = round(timed(lambda: eng.select_victims(s.snapshot, s.pending, 0, None, pick=True), 20), 1)
print(json.dumps(out))

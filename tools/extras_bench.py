"""Cost of the optional stages on C2 (5 000 nodes x 100 000 pods), one B200: node sampling, nominated pods, reason
planes — wall time of fgs_evaluate with and without each. python tools/extras_bench.py"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import flex_gpu_scheduler_b200 as fgs  # noqa: E402
from flex_gpu_scheduler_b200 import _lib as L  # noqa: E402

s = fgs.config_synth(2, n_nominated=64)
eng = fgs.Engine(device=0)
eng.set_plugins(fgs.config_plugins(2))
eng.upload_nodes(fgs.Packer(2).nodes(s.snapshot))
t = fgs.Packer(2).pods(s.pending)


def timed(f, reps=5):
    f()
    t0 = time.perf_counter()
    for _ in range(reps):
        r = f()
    return 1e3 * (time.perf_counter() - t0) / reps, r


base, r0 = timed(lambda: eng.evaluate(t))
samp, r1 = timed(lambda: eng.evaluate(t, sample=1, start_node_index=0))
Q = int(s.nominated.contents.n_pods)
nodes = np.random.default_rng(3).integers(0, s.n_nodes, Q).astype(np.int32)
eng.set_nominated(fgs.Packer(2).pods(s.nominated), nodes)
nom, r2 = timed(lambda: eng.evaluate(t))
eng.set_nominated()
small = fgs.Packer(2).pods(s.pending, 0, 4096)
eng.evaluate(small)
reason, _ = timed(lambda: eng.reason_plane(L.FGS_PL_FIT), reps=3)
print(json.dumps({"evaluate_ms": base, "evaluate_sampled_ms": samp, "sampling_extra_ms": samp - base,
                  "mean_feasible_all": float(r0.feasible_count.mean()), "mean_feasible_sampled": float(r1.feasible_count.mean()),
                  "evaluate_with_64_nominated_pods_ms": nom, "nominated_extra_ms": nom - base,
                  "pairs_lost_to_nominated": int(r0.feasible_count.sum() - r2.feasible_count.sum()),
                  "fit_reason_plane_4096_pods_ms_incl_d2h": reason}))

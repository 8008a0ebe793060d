"""Node sampling on a long batch: fgs_evaluate(sample=1) of 100 000 pods on 5 000 nodes with the walk composed per block of
pods and with the serial chain (FGS_SAMPLE_SERIAL=1), and the same evaluation without sampling."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import flex_gpu_scheduler_b200 as fgs  # noqa: E402

n_nodes, n_pods = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (5000, 100000)
s = fgs.config_synth(2, n_nodes=n_nodes, n_pending=n_pods)
cfg = fgs.config_plugins(2)
t = fgs.Packer(2).pods(s.pending)
for label, env in (("composed", None), ("serial", "1")):
    if env:
        os.environ["FGS_SAMPLE_SERIAL"] = env
    eng = fgs.Engine(device=0)
    eng.set_plugins(cfg)
    eng.upload_nodes(fgs.Packer(2).nodes(s.snapshot))
    eng.evaluate(t)
    t0 = time.perf_counter()
    for _ in range(3):
        eng.evaluate(t)
    base = (time.perf_counter() - t0) / 3
    eng.evaluate(t, sample=1)
    t0 = time.perf_counter()
    for _ in range(3):
        r = eng.evaluate(t, sample=1, start_node_index=17)
    samp = (time.perf_counter() - t0) / 3
    print(f"{n_nodes} nodes x {n_pods} pods, {label} walk: evaluate {base * 1e3:.2f} ms, with sampling {samp * 1e3:.2f} ms, next start {r.next_start_node_index}")
    eng.close()
    os.environ.pop("FGS_SAMPLE_SERIAL", None)

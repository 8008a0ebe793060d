"""Sequential scheduling cycles per second on one GPU (fgs_schedule_sequential): python tools/seq_bench.py [cfg] [nodes] [pods]"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import flex_gpu_scheduler_b200 as fgs  # noqa: E402

cfg_id = int(sys.argv[1]) if len(sys.argv) > 1 else 2
n_nodes = int(sys.argv[2]) if len(sys.argv) > 2 else 5000
n_pods = int(sys.argv[3]) if len(sys.argv) > 3 else 100000
s = fgs.config_synth(cfg_id, n_nodes=n_nodes, n_pending=n_pods)
cfg = fgs.config_plugins(cfg_id)
pk_n, pk_p = fgs.Packer(2), fgs.Packer(2)
nodes = pk_n.nodes(s.snapshot)
table = pk_p.pods(s.pending)
for ctas in [None] + [int(x) for x in os.environ.get("SEQ_SWEEP", "").split(",") if x]:
    if ctas is not None:  # 0 = automatic geometry without the cluster variant
        if ctas == 0:
            os.environ["FGS_SEQ_CLUSTER"] = "0"
        else:
            os.environ["FGS_SEQ_CTAS"] = str(ctas)
    eng = fgs.Engine(device=0)
    eng.set_plugins(cfg)
    for rep in range(2):
        eng.upload_nodes(nodes)
        t0 = time.perf_counter()
        res = eng.schedule_sequential(table)
        wall = time.perf_counter() - t0
    t = eng.timings()
    print(json.dumps(dict(cfg=cfg_id, nodes=n_nodes, pods=n_pods, ctas=t["seq_ctas"], kernel_ms=t["seq_ms"],
                          cycles_per_s=n_pods / (t["seq_ms"] * 1e-3), us_per_cycle=t["seq_ms"] * 1e3 / n_pods, wall_ms=wall * 1e3,
                          committed=int(res.committed.sum()), unschedulable=int((res.feasible_count == 0).sum()),
                          errors=int((res.pod_err != 0).sum()))))
    eng.close()

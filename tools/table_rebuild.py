import os, sys, time
sys.path.insert(0, "/root/repo")
import numpy as np
import flex_gpu_scheduler_b200 as fgs
for n in (1000, 5000, 16000, 50000):
    s = fgs.config_synth(2, n_nodes=n, n_pending=2000)
    cfg = fgs.config_plugins(2)
    eng = fgs.Engine(device=0)
    eng.set_plugins(cfg)
    pk = fgs.Packer(2)
    eng.upload_nodes(pk.nodes(s.snapshot))
    t = fgs.Packer(2).pods(s.pending)
    eng.evaluate(t)
    ts = []
    for k in range(6):
        eng.update_nodes(fgs.Packer(2).nodes(s.snapshot, 7, 8), [7])   # node table changes: tables stale
        t0 = time.perf_counter()
        eng.evaluate(t)                                               # 2000 pods: rebuild + evaluate
        ts.append(time.perf_counter() - t0)
    eng.evaluate(t)
    t0 = time.perf_counter(); eng.evaluate(t); t1 = time.perf_counter() - t0
    print(f"nodes {n}: evaluate after a node change {sorted(ts)[len(ts)//2]*1e6:.0f} us, with current tables {t1*1e6:.0f} us ({os.environ.get('FGS_FT_SORT','smem')})")
    eng.close()

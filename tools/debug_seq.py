import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import flex_gpu_scheduler_b200 as fgs
s = fgs.config_synth(1, n_nodes=700, n_pending=260, max_existing_pods=6, pm_panic_nodes=10, pm_invalid_nodes=10, pm_nongpu_nodes=50)
cfg = fgs.config_plugins(1)
pk = fgs.Packer(2)
one = fgs.Packer(2).pods(s.pending, 0, 1)
eng = fgs.Engine(device=0); eng.set_plugins(cfg); eng.upload_nodes(pk.nodes(s.snapshot))
b = eng.evaluate(one, want_planes=True)
feas = b.feasible(700)[0]
print("batch count", int(b.feasible_count[0]), int(feas.sum()))
bad = []
for n in range(700):
    e2 = fgs.Engine(device=0); e2.set_plugins(cfg); e2.upload_nodes(pk.nodes(s.snapshot, n, n + 1))
    r = e2.schedule_sequential(one)
    if int(r.feasible_count[0]) != int(feas[n]):
        bad.append(n)
    e2.close()
print("bad nodes", bad)
t = pk.nodes(s.snapshot).contents
for n in bad[:6]:
    print(n, "flags", t.flags[n], "mono", t.monopoly_mask[n], "alloc_pods", t.alloc_pods[n], "num_pods", t.num_pods[n],
          "gpu", t.alloc_scalar[n], "mem", t.alloc_scalar[700 + n], "gsum", t.flex_gpu_sum[n], "msum", t.flex_mem_sum[n],
          "used", [t.used_mem[d * 700 + n] for d in range(t.n_devices)], "req", t.req_mcpu[n], t.req_mem[n], "alloc", t.alloc_mcpu[n], t.alloc_mem[n])
pt = one.contents
print("pod", pt.flex_gpu_limit[0], pt.flex_mem_limit[0], pt.flags[0], pt.req_mcpu[0], pt.req_mem[0], pt.req_eph[0], pt.req_scalar[0], pt.req_scalar[1], pt.req_scalar_present[0])

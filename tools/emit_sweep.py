"""Emit / Filter pass times on ONE GPU for a given shard shape, e.g. the shard a GPU holds when config 5 is sharded over 8
(6 250 nodes x 65 536 pods): python tools/emit_sweep.py 6250 65536. The launch geometry is read from the environment
(FGS_WAVES_EMIT, …) once per process, so sweep it from the shell."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import flex_gpu_scheduler_b200 as fgs  # noqa: E402

n_nodes, n_pods = int(sys.argv[1]), int(sys.argv[2])
cfg_id = int(sys.argv[3]) if len(sys.argv) > 3 else 2
s = fgs.config_synth(cfg_id, n_nodes=n_nodes, n_pending=n_pods)
stream = torch.cuda.Stream()
e = fgs.Engine(device=0, stream=stream.cuda_stream)
e.set_plugins(fgs.config_plugins(cfg_id))
e.upload_nodes(fgs.Packer(2).nodes(s.snapshot))
e.stage_pods(fgs.Packer(2).pods(s.pending))
for _ in range(3):
    e.pass_reduce()
    e.pass_emit()
e.synchronize()
K = 10
ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(K)]
for k in range(K):
    ev[k][0].record(stream)
    e.pass_reduce()
    ev[k][1].record(stream)
    e.pass_emit()
    ev[k][2].record(stream)
torch.cuda.synchronize()
red = sorted(a.elapsed_time(b) for a, b, _ in ev)[K // 2]
emit = sorted(b.elapsed_time(c) for _, b, c in ev)[K // 2]
gb = n_pods * n_nodes * 8 / 1e9
print(f"nodes {n_nodes} pods {n_pods} waves_emit={os.environ.get('FGS_WAVES_EMIT', 'default')}: filter {red * 1e3:.1f} us, emit {emit * 1e3:.1f} us "
      f"= {gb / (emit * 1e-3):.0f} GB/s")

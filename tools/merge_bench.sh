#!/bin/bash
# tools/merge_bench.sh N — the node-sharded bench with each kind of cross-shard merge, one JSON line each in gpurun_out/
N=${1:-2}
mkdir -p gpurun_out
for m in ${MODES:-auto peer nccl}; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 200)) \
    bench.py --gpus $N --steps 20 --warmup 3 --merge $m > gpurun_out/merge_${N}_$m.json 2> gpurun_out/merge_${N}_$m.err
  echo "rc=$? merge=$m"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/merge_${N}_$m.json").read().strip().splitlines()[-1])
    print(d["n_gpus"], round(d["ms_per_step"], 4), "%.3e" % d["value"], {k: round(v, 4) for k, v in d["kernel_ms"].items() if not isinstance(v, str)}, d["config"]["merge"], d["e2e"]["value"])
except Exception as ex:
    print("no line:", ex)
PY
done

#!/bin/bash
# tools/peer_ctas_sweep.sh N — grid size of the peer-memory merge kernels against the overlapped step time
N=${1:-2}
mkdir -p gpurun_out
for c in ${CTAS:-8 16 32}; do
  FGS_PEER_CTAS_OVERLAP=$c timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 200)) \
    bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ctas_${N}_$c.json 2> gpurun_out/ctas_${N}_$c.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/ctas_${N}_$c.json").read().strip().splitlines()[-1])
    print("ctas $c", d["n_gpus"], round(d["ms_per_step"], 4), "%.3e" % d["value"], round(d["kernel_ms"].get("one_batch_at_a_time_ms_per_step", 0), 4), "e2e %.3e" % d["e2e"]["value"])
except Exception as ex:
    print("no line:", ex)
PY
done

for wr in 2 3 4; do for we in 2 3 4 6 8; do
FGS_WAVES_REDUCE=$wr FGS_WAVES_EMIT=$we python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('wr=$wr we=$we', d['kernel_ms'], round(d['value']/1e9,1))"
done; done

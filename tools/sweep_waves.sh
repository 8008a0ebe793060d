# CTAs per resident slot of the two passes (FGS_WAVES_REDUCE / FGS_WAVES_EMIT) against the kernel times on C2
for wr in ${WR:-3}; do for we in ${WE:-1 2 3 4 6 8}; do
FGS_WAVES_REDUCE=$wr FGS_WAVES_EMIT=$we python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-sequential 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('wr=$wr we=$we', {k: round(v, 4) for k, v in d['kernel_ms'].items()}, round(d['value']/1e9,1))"
done; done

mkdir -p gpurun_out
make -C benchmark_dolfinx_b200 -j 16 2>&1 | tail -1
one() { BD_SLAB_TUNE=$1 timeout 120 python bench.py --degree $2 --ndofs $3 --float $4 --qmode $5 --no-cpu --no-e2e --no-cg --steps 10 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print('tune $1 Q$2 f$4 q$5', 'value',round(d['value'],2),'frac',round(r['frac'],3),'kernel_ms',round(r['kernel_ms_per_apply'],2))
"; }
for v in 0 2 3 4; do one $v 4 300000000 64 1; done
for v in 0 2 3 4; do one $v 5 300000000 64 0; done
for v in 0 2 3; do one $v 3 500000000 32 1; done

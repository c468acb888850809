mkdir -p gpurun_out
make -C benchmark_dolfinx_b200 -j 16 2>&1 | tail -1
for w in 2 4; do for pf in 0 1; do for cfg in "6 500000000 1" "7 300000000 0"; do
set -- $cfg
BD_MMA8_WARPS=$w BD_MMA8_PREFETCH=$pf python bench.py --degree $1 --ndofs $2 --qmode $3 --no-cpu --no-e2e --no-cg --steps 10 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print('warps $w prefetch $pf Q$1 q$3', 'value',round(d['value'],2),'frac',round(r['frac'],3),'kernel_ms',round(r['kernel_ms_per_apply'],2))
"
done; done; done

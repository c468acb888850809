mkdir -p gpurun_out
make -C benchmark_dolfinx_b200 -j 16 2>&1 | tail -1
for v in 1 6 7 8 1 0; do
BD_SLAB_TUNE=$v python bench.py --no-cpu --no-e2e --no-cg --steps 10 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print('tune $v', 'value',round(d['value'],2),'frac',round(r['frac'],3),'kernel_ms',round(r['kernel_ms_per_apply'],2))
"
done

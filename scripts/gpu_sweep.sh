# usage: TAG=v7 CFGS="3:300000000:64:1 6:500000000:64:1" bash scripts/gpu_sweep.sh
mkdir -p gpurun_out
make -C benchmark_dolfinx_b200 -j 16 2>&1 | tail -1
if [ -n "$TESTS" ]; then python -m pytest $TESTS -x -q -m gpu 2>&1 | tail -4; fi
for cfg in $CFGS; do
IFS=: read d n f q <<< "$cfg"
python bench.py --degree $d --ndofs $n --float $f --qmode $q --no-cpu --no-e2e $EXTRA > gpurun_out/bench_${TAG}_q${d}_f${f}_m${q}.json 2>&1
python - gpurun_out/bench_${TAG}_q${d}_f${f}_m${q}.json <<'PY'
import json,sys
ok=False
for l in open(sys.argv[1]):
    if l.startswith('{'):
        ok=True
        d=json.loads(l); r=d['roofline']; print(sys.argv[1].split('/')[-1], 'value',round(d['value'],2),'frac',round(r['frac'],3),'kernel_ms',round(r['kernel_ms_per_apply'],2),'cg',d['cg'] and round(d['cg']['value'],2), 'gflops', round(r['fp_gflops_algorithmic']))
if not ok: print(open(sys.argv[1]).read()[-1200:])
PY
done

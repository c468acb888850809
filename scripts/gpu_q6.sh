set -x
mkdir -p gpurun_out
make -C benchmark_dolfinx_b200 -j 16 2>&1 | tail -1
python -m pytest tests/test_gpu_operator.py tests/test_gpu_cg.py -x -q -m gpu 2>&1 | tail -3
for cfg in "6 500000000 64 1" "7 300000000 64 0"; do
set -- $cfg
python bench.py --degree $1 --ndofs $2 --float $3 --qmode $4 --no-cpu --no-e2e > gpurun_out/bench_v6_q$1_f$3_m$4.json 2>&1; echo "rc=$?"
python - gpurun_out/bench_v6_q$1_f$3_m$4.json <<'PY'
import json,sys
ok=False
for l in open(sys.argv[1]):
    if l.startswith('{'):
        ok=True
        d=json.loads(l); r=d['roofline']; print(sys.argv[1], 'value',round(d['value'],2),'frac',round(r['frac'],3),'kernel_ms',round(r['kernel_ms_per_apply'],2),'cg',d['cg'] and round(d['cg']['value'],2), 'gflops', round(r['fp_gflops_algorithmic']))
if not ok: print(open(sys.argv[1]).read()[-1500:])
PY
done
